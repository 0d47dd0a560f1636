/*
 * msda_b200.h -- C ABI of libmsda_b200.so: multi-scale deformable attention
 * (forward + backward) for NVIDIA B200 (sm_100a).
 *
 * This is the drop-in boundary for the reference's native module
 * `MultiScaleDeformableAttention` (pybind, models/ops/src/vision.cpp:13-16):
 *
 *   ms_deform_attn_forward (models/ops/src/ms_deform_attn.h:20-39,
 *                           models/ops/src/cuda/ms_deform_attn_cuda.cu:20-80)
 *       -> msda_forward
 *   ms_deform_attn_backward(models/ops/src/ms_deform_attn.h:41-61,
 *                           models/ops/src/cuda/ms_deform_attn_cuda.cu:83-153)
 *       -> msda_backward (+ msda_backward_workspace_bytes)
 *
 * Conventions
 *   - plain pointers and sizes; no torch / pybind types.
 *   - every data pointer is a DEVICE pointer; the library borrows it for the
 *     duration of the enqueue only and never allocates or retains memory.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - work is enqueued asynchronously; no host synchronisation is performed.
 *   - return value: MSDA_OK (0) or a negative msda_status; never throws/exits.
 *     (The reference only printf()s launch errors, .cuh:949-953 -- this ABI is
 *     intentionally stricter.)
 *   - no static mutable state: re-entrant, callable from several host threads
 *     on different streams.
 *
 * Tensor layouts (row-major contiguous, identical to the reference):
 *   value              [N, S, M, D]          S = sum_l H_l*W_l
 *   spatial_shapes     [L, 2] int64 (H_l, W_l)          -- device memory
 *   level_start_index  [L]    int64                     -- device memory
 *   sampling_loc       [N, Lq, M, L, P, 2]   normalised (x -> W, y -> H)
 *   attn_weight        [N, Lq, M, L, P]
 *   output/grad_output [N, Lq, M*D]
 */
#ifndef MSDA_B200_H_
#define MSDA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSDA_B200_ABI_VERSION 2

/* element type of value / sampling_loc / attn_weight / output / all grads */
typedef enum {
  MSDA_F32 = 0,  /* reference: AT_DISPATCH_FLOATING_TYPES (cuda.cu:64) */
  MSDA_F64 = 1,  /* reference: same macro; needed by gradcheck (test.py:76) */
  MSDA_BF16 = 2  /* new capability (fp32 accumulate) */
} msda_dtype;

/* handling of taps outside the feature map */
typedef enum {
  MSDA_PAD_ZEROS = 0, /* reference CUDA op (.cuh:56-79, gate at :289) */
  MSDA_PAD_BORDER = 1 /* reference PyTorch path (grid_sample border,
                         functions/ms_deform_attn_func.py:61-62) */
} msda_padding;

typedef enum {
  MSDA_OK = 0,
  MSDA_ERR_NULL_POINTER = -1,
  MSDA_ERR_BAD_SHAPE = -2,      /* a dimension <= 0 (N, Lq, S may be 0) */
  MSDA_ERR_BAD_DTYPE = -3,
  MSDA_ERR_BAD_PADDING = -4,
  MSDA_ERR_TOO_LARGE = -5,      /* exceeds the 2^31 row / group index range */
  MSDA_ERR_WORKSPACE = -6,      /* workspace missing or too small */
  MSDA_ERR_CUDA = -7,           /* CUDA runtime error (see msda_last_cuda_error) */
  MSDA_ERR_MISALIGNED = -8,     /* a data pointer is not 16-byte aligned */
  MSDA_ERR_UNSUPPORTED = -10,   /* msda_module_*: the call is outside the fused path (nothing was
                                   enqueued); use msda_frontend_* + msda_forward_ex / msda_backward_ex */
  MSDA_ERR_SCRATCH_LIMIT = -11, /* msda_forward_static: the call needs more scratch than the
                                   library's built-in MSDA_STATIC_SCRATCH_BYTES (nothing was
                                   enqueued); use msda_forward_ex with a caller workspace */
  MSDA_ERR_NO_TMA = -9          /* internal: the TMA descriptor could not be encoded; the
                                   entry points fall back to the row-window kernels and
                                   never return it */
} msda_status;

/* flags for msda_forward / msda_backward (bitwise or) */
#define MSDA_FLAG_FORCE_GENERIC 1u /* use the any-D scalar kernels even when the
                                      vectorised path applies (testing aid) */
#define MSDA_FLAG_SKIP_ZERO_FILL 2u /* msda_backward: the caller has already
                                      zeroed the accumulation target (grad_value
                                      for f32/f64, the workspace for bf16) */
#define MSDA_FLAG_NO_WINDOW 4u      /* never use the row-window kernels
                                      (msda_win.cuh); per-tap kernels only */
#define MSDA_FLAG_FORCE_WINDOW 8u   /* use them whenever they are instantiated for
                                      the shape, also for short query lists
                                      (default: Lq >= 256) */
#define MSDA_FLAG_DETERMINISTIC 16u /* msda_backward: bitwise reproducible grad_value.
                                      Contributions are accumulated as 64-bit fixed
                                      point with integer atomics (order-independent)
                                      by the kernel family the shape selects (tile,
                                      row-window, per-tap or any-D), then converted;
                                      needs the workspace
                                      msda_backward_workspace_bytes_flags reports
                                      (32 B + 8 B per grad_value element).  A
                                      non-finite grad_output / attn_weight makes the
                                      whole grad_value NaN.  Slower: an optional
                                      variant, as the sorted scatter of the design
                                      brief, without the sort */
/* Semantics of the default kernels on NON-FINITE value rows (temporal maps, >= 256 queries: CTA-tile and
 * row-window kernels): a block of 4 time-adjacent queries multiplies every row of its window by each of
 * its queries' weights, so an Inf / NaN value row reaches the block's other queries as 0 * Inf = NaN,
 * where the reference poisons only the queries that tap the row -- a superset of the reference's NaN
 * outputs, never outside the tapping blocks; finite outputs are equal.  MSDA_FLAG_NO_WINDOW selects the
 * per-tap kernels, which match the reference exactly. */
#define MSDA_FLAG_NO_TILE 32u       /* never use the CTA-tile kernels (msda_tile.cuh:
                                      time-sorted query tiles, TMA-filled shared
                                      memory); the row-window kernels run instead */
#define MSDA_FLAG_PLAN 64u          /* msda_forward_ex: run the CTA-tile path as a PLAN kernel
                                      (geometry, windows and merged weights of every tile
                                      -> 13 KB per tile in the workspace) and a persistent
                                      APPLY kernel (TMA producer warp + consumer warps),
                                      msda_plan.cuh.  Measured slower than the one-kernel
                                      form on B200: kept as an
                                      experiment, off by default */
#define MSDA_FLAG_SAVE_PLAN 128u    /* msda_forward_ex + msda_backward_ex, CTA-tile shape class: the
                                      forward also SAVES every tile's plan (windows + merged
                                      weights, ~14.6 KB reserved / ~7 KB written per 64 queries
                                      and head) behind the query order in its workspace
                                      (msda_forward_workspace_bytes with this flag reports the
                                      size; 128-byte aligned); msda_backward_ex given that
                                      workspace as `plan` AND this flag starts every tile from
                                      the saved plan instead of the sampling points.  Set it on
                                      both calls or on neither; a forward workspace too small for
                                      the plans just keeps the query order (then do not set the
                                      flag on the backward).  Same results.  An EXPERIMENT
                                      (measured: no gain): honoured only by a
                                      library built with -DMSDA_TILE_SAVED_PLANS=1; the default build
                                      ignores the flag on both calls (msda_forward_workspace_bytes
                                      then reports the same size with and without it) */
#define MSDA_FLAG_PERSIST (1u << 20) /* msda_forward_ex, fp32 CTA-tile shape class: the forward tile kernel
                                      in its persistent form (one CTA per resident slot walks tiles and
                                      holds the next tile's query list and points ahead of time).  Same
                                      results; an experiment */
/* bits 8..19: queries-per-tile multiplier override for the vectorised kernels
 * (0 = heuristic); a tuning knob, results do not depend on it. */
#define MSDA_FLAG_QTILE(k) (((unsigned)(k) & 0xfffu) << 8)

/* Library / ABI identification. */
int msda_abi_version(void);
const char* msda_error_string(int status);
/* Name of the kernel family msda_forward/backward would pick:
 * "fast" (vectorised, warp-cooperative) or "generic". */
const char* msda_kernel_family(int64_t D, int dtype, unsigned flags);
/* Kernel family for a full problem shape (16-byte aligned pointers assumed, and for
 * "tile" a plan workspace: msda_forward_ex / msda_backward_ex):
 * "tile"    CTA-tile kernels msda_{fwd,bwd}_tile_kernel (the same shape class as
 *           "window"): 64 time-adjacent queries per CTA -- the queries are taken
 *           in time order, whatever their index order --, the value rows they tap
 *           brought to shared memory by TMA; a tile whose rows do not fit runs the
 *           per-tap body,
 * "window"  row-window kernels msda_{fwd,bwd}_win_kernel (temporal maps, P = 4,
 *           L <= 4, long query lists; per block of 4 queries they fall back to
 *           the per-tap body when the block has no compact window),
 * "fast"    per-tap vectorised kernels msda_{fwd,bwd}_fast_kernel,
 * "generic" any-D / fp64 kernels, "invalid" for a bad shape or dtype. */
const char* msda_kernel_choice(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                               int64_t Lq, int64_t P, int dtype, unsigned flags);

/*
 * out[b,q,m,:] = sum_{l,p} aw[b,q,m,l,p] * bilinear(value_l[b,:,m,:], loc[b,q,m,l,p])
 * Replaces ms_deform_attn_cuda_forward (cuda.cu:20-80) + the im2col kernel
 * (.cuh:238-300).  `output` is fully written (no pre-zeroing needed).
 * The reference's im2col_step batch chunking is a host-side knob validated by
 * the Python shim; results do not depend on it, so it is not part of the ABI.
 */
int msda_forward(const void* value, const int64_t* spatial_shapes,
                 const int64_t* level_start_index, const void* sampling_loc,
                 const void* attn_weight, void* output,
                 int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                 int64_t Lq, int64_t P,
                 int dtype, int padding, unsigned flags, void* stream);

/*
 * The same with a caller-provided workspace of msda_forward_workspace_bytes() bytes
 * (16-byte aligned device memory, borrowed for the call).  For the "tile" shape class
 * the call first writes the PLAN there -- the time order of each batch item's queries,
 * int32 [N][Lq] -- and runs the CTA-tile kernels; the same buffer, untouched, can be
 * handed to msda_backward_ex as `plan` so that the backward does not sort again.
 * Without a workspace (or with one that is too small) the call is msda_forward.
 */
size_t msda_forward_workspace_bytes(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                                    int64_t Lq, int64_t P, int dtype, unsigned flags);
int msda_forward_ex(const void* value, const int64_t* spatial_shapes,
                    const int64_t* level_start_index, const void* sampling_loc,
                    const void* attn_weight, void* output,
                    int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                    int64_t Lq, int64_t P,
                    int dtype, int padding, unsigned flags,
                    void* workspace, size_t workspace_bytes, void* stream);

/*
 * msda_forward_ex for callers that cannot allocate (code under CUDA graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * The plan of the CTA-tile kernels (the
 * time order of the queries, int32 [N][Lq]) goes to a scratch array built into the
 * library (a file-scope __device__ array of MSDA_STATIC_SCRATCH_BYTES, resident on
 * every GPU that loads the library until the process exits) instead of a caller
 * workspace; the call itself allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  Same kernels, same results as msda_forward_ex with a
 * workspace of msda_forward_workspace_bytes().
 *   - msda_forward_workspace_bytes(..., flags without MSDA_FLAG_PLAN) >
 *     MSDA_STATIC_SCRATCH_BYTES (N * Lq > 262,144 queries) -> MSDA_ERR_SCRATCH_LIMIT,
 *     nothing enqueued.  MSDA_FLAG_PLAN is ignored (its plan is 13 KB per tile).
 *   - a shape class that needs no scratch (per-tap / row-window / any-D kernels) runs
 *     as msda_forward, which never allocated.
 *   - ALL calls on a device share the array: two calls that overlap on the GPU
 *     (different streams, or different host threads on different streams) race on it.
 *     Calls on one stream are ordered and safe.  The array keeps the last call's
 *     plan; every call overwrites the [N][Lq] entries it reads before reading them,
 *     so nothing in it has to start at zero.
 */
#define MSDA_STATIC_SCRATCH_BYTES (1u << 20)
int msda_forward_static(const void* value, const int64_t* spatial_shapes,
                        const int64_t* level_start_index, const void* sampling_loc,
                        const void* attn_weight, void* output,
                        int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                        int64_t Lq, int64_t P,
                        int dtype, int padding, unsigned flags, void* stream);

/* Bytes of scratch msda_backward takes.  REQUIRED part: nothing for f32/f64, an fp32
 * accumulation image of grad_value for bf16 (the deterministic variant: see the
 * _flags form).  OPTIONAL part, for the "tile" shape class: room for the plan behind
 * it (256-byte aligned); a workspace without that room selects the row-window
 * kernels. */
size_t msda_backward_workspace_bytes(int64_t N, int64_t S, int64_t M, int64_t D,
                                     int64_t L, int64_t Lq, int64_t P, int dtype);

/* The same for a given `flags` word (MSDA_FLAG_DETERMINISTIC changes the answer). */
size_t msda_backward_workspace_bytes_flags(int64_t N, int64_t S, int64_t M, int64_t D,
                                           int64_t L, int64_t Lq, int64_t P, int dtype,
                                           unsigned flags);

/*
 * Replaces ms_deform_attn_cuda_backward (cuda.cu:83-153) + the col2im kernels
 * (.cuh:302-921).  All three gradient buffers are fully defined on return
 * (grad_value is zero-filled inside the call on `stream`; callers need not
 * pre-zero anything).
 */
int msda_backward(const void* value, const int64_t* spatial_shapes,
                  const int64_t* level_start_index, const void* sampling_loc,
                  const void* attn_weight, const void* grad_output,
                  void* grad_value, void* grad_sampling_loc, void* grad_attn_weight,
                  int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                  int64_t Lq, int64_t P,
                  int dtype, int padding, unsigned flags,
                  void* workspace, size_t workspace_bytes, void* stream);

/* msda_backward with the plan msda_forward_ex left in its workspace for the SAME
 * sampling_loc (NULL: the plan is computed into the workspace tail, as msda_backward
 * does).  What MSDeformAttnFunction.backward calls (ms_deform_attn_func.py:33-41):
 * autograd keeps the forward's plan next to the saved tensors. */
int msda_backward_ex(const void* value, const int64_t* spatial_shapes,
                     const int64_t* level_start_index, const void* sampling_loc,
                     const void* attn_weight, const void* grad_output,
                     void* grad_value, void* grad_sampling_loc, void* grad_attn_weight,
                     int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                     int64_t Lq, int64_t P,
                     int dtype, int padding, unsigned flags,
                     void* workspace, size_t workspace_bytes, const void* plan, void* stream);

/*
 * Module front-end: what MSDeformAttn.forward does between its Linear layers and
 * the op (models/ops/modules/ms_deform_attn.py:98-117; live twin
 * models/modules/attention.py:462-496) -- softmax of the attention logits over
 * L*P, sampling locations from reference points and offsets, the constant
 * y = 0.5 -- in one kernel instead of a softmax and four or five elementwise
 * kernels.  All tensors contiguous, same dtype:
 *   offsets [N,Lq,M,L,P]   logits [N,Lq,M,L*P]   ref [N,Lq,L,ref_dim]
 *   level_sizes [L] int64 (device): T_l, the divisor of the offsets (ref_dim 1)
 *   loc [N,Lq,M,L,P,2] = (ref0 + off / T_l, 0.5)                       ref_dim 1
 *                        (ref0 + off / P * ref1 * 0.5, 0.5)            ref_dim 2
 *   aw  [N,Lq,M,L,P]   = softmax(logits)
 * SURVEY.md section 8(f) row 1 ("next"): first step, the arithmetic fused into one
 * streaming kernel; folding it into the gather kernels themselves is later work.
 */
int msda_frontend_forward(const void* offsets, const void* logits, const void* ref,
                          const int64_t* level_sizes, void* loc, void* aw,
                          int64_t N, int64_t Lq, int64_t M, int64_t L, int64_t P,
                          int ref_dim, int dtype, void* stream);

/*
 * Gradients of the above: grad_offsets, grad_logits (softmax backward from the
 * saved aw) and, when grad_ref is not NULL, grad_ref [N,Lq,L,ref_dim] (summed
 * over heads and points; fully written, no pre-zeroing needed).  The y half of
 * grad_loc is ignored (y is a constant).
 */
int msda_frontend_backward(const void* offsets, const void* aw, const void* ref,
                           const int64_t* level_sizes, const void* grad_loc,
                           const void* grad_aw, void* grad_offsets, void* grad_logits,
                           void* grad_ref,
                           int64_t N, int64_t Lq, int64_t M, int64_t L, int64_t P,
                           int ref_dim, int dtype, void* stream);

/*
 * The module's front-end INSIDE the op (second step): for the
 * shape class of the CTA-tile kernels (temporal maps (1, T_l), P = 4, L <= 4, D = 64 (f32, bf16) or
 * 32 (f32), Lq >= 256) the tile kernels take what MSDeformAttn.forward holds BEFORE
 * models/ops/modules/ms_deform_attn.py:98-117 -- offsets [N,Lq,M,L,P], logits [N,Lq,M,L*P],
 * ref [N,Lq,L,ref_dim] -- and do the softmax, x = ref + off / T_l (ref_dim 1) or
 * ref0 + off / P * ref1 * 0.5 (ref_dim 2) and y = 0.5 in their prologue; sampling_locations and
 * attention_weights are never written to memory (37 MB at the encoder shape, and two kernels per
 * module call).  The backward writes grad_offsets, grad_logits and (fp32, summed over heads and
 * points, zero-filled inside; may be NULL) grad_ref directly.  `workspace` of the forward:
 * msda_forward_workspace_bytes(..., flags) bytes, which is the `plan` the backward needs;
 * `workspace` of the backward: msda_backward_workspace_bytes (bf16 image) or NULL for f32.
 * padding_mask [N, S] bool (true = padding; may be NULL): value.masked_fill(mask, 0) of
 * ms_deform_attn.py:96-97 without the pass over value -- masked rows are zeroed in the shared-memory
 * tile as they land, and receive no gradient.
 * msda_module_supported() != 0 <=> the pair accepts the shape; otherwise they return
 * MSDA_ERR_UNSUPPORTED without enqueueing anything.  spatial_shapes must hold H_l = 1.
 */
int msda_module_supported(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq,
                          int64_t P, int ref_dim, int dtype, unsigned flags);
int msda_module_forward(const void* value, const int64_t* spatial_shapes,
                        const int64_t* level_start_index, const void* offsets, const void* logits,
                        const void* ref, const unsigned char* padding_mask, void* output,
                        int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P,
                        int ref_dim, int dtype, int padding, unsigned flags,
                        void* workspace, size_t workspace_bytes, void* stream);
int msda_module_backward(const void* value, const int64_t* spatial_shapes,
                         const int64_t* level_start_index, const void* offsets, const void* logits,
                         const void* ref, const unsigned char* padding_mask,
                         const void* grad_output, void* grad_value,
                         void* grad_offsets, void* grad_logits, float* grad_ref,
                         int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P,
                         int ref_dim, int dtype, int padding, unsigned flags,
                         void* workspace, size_t workspace_bytes, const void* plan, void* stream);

/*
 * Decoder attention map (utils/dam.py:20-73 attn_map_to_flat_grid; consumer
 * models/criterion.py:246-283): the attention weight of every sampling point
 * scattered onto the flat token grid, summed over queries and points.
 *   loc [BL,Q,M,L,P,loc_cols] (x in column 0; loc_cols 1 or 2)   aw [BL,Q,M,L,P]
 *   level_sizes, level_start [L] int64 (device)
 *   out [BL,M,S] float32, zero-filled inside the call
 * out[bl,m,start_l + floor(x*T_l)] += aw * (frac - 1),  [.. + 1] += aw * frac,
 * indices outside [0, T_l) dropped -- the reference's arithmetic, sign included.
 * BL = batch * decoder layers.  SURVEY.md section 8(f) row 2.
 */
int msda_dam_scatter(const void* loc, const void* aw, const int64_t* level_sizes,
                     const int64_t* level_start, float* out,
                     int64_t BL, int64_t Q, int64_t M, int64_t L, int64_t P, int64_t S,
                     int loc_cols, int dtype, void* stream);

/*
 * Encoder reference points (DeformableTransformerEncoder.get_reference_points,
 * models/deformable/unimodal_deformable_transformer.py:259-276 and its copies in
 * the multimodal and sparse encoders):
 *   out[n, s, l] = (t + 0.5) / (valid_ratios[n, k] * T_k) * valid_ratios[n, l]
 * for token s = level_start_k + t.  level_sizes [L] int64 and valid_ratios [N,L]
 * fp32 on the device; out [N,S,L] fp32 (the reference returns it as [N,S,L,1]).
 * SURVEY.md section 8(f) row 4.
 */
int msda_encoder_reference_points(const int64_t* level_sizes, const float* valid_ratios,
                                  float* out, int64_t N, int64_t S, int64_t L, void* stream);

/*
 * prepare_encoder_inputs (models/deformable/unimodal_deformable_transformer.py:90-134 and its copies
 * in the multimodal and sparse transformers; SURVEY.md section 8(f) row 4, second half): the per-level
 * transposes, the level-embedding add, the three concatenations and the valid ratios in two launches.
 *   srcs[l], pos_embeds[l]  [B, C, T_l]  (HOST arrays of L device pointers)    masks[l] [B, T_l] bool
 *   level_sizes_host [L] int64 in HOST memory (the T_l are tensor shapes: no device round trip)
 *   level_embed [L, C]
 *   -> src_flatten, lvl_pos_embed_flatten [B, S, C] (pos + level_embed[l]), mask_flatten [B, S] bool,
 *      valid_ratios [B, L] float32 = (tokens of the level that are not padding) / T_l
 */
int msda_prepare_encoder_inputs(const void* const* srcs, const void* const* pos_embeds,
                                const unsigned char* const* masks, const int64_t* level_sizes_host,
                                const void* level_embed, void* src_flatten,
                                void* lvl_pos_embed_flatten, unsigned char* mask_flatten,
                                float* valid_ratios, int64_t B, int64_t C, int64_t L, int dtype,
                                void* stream);

/*
 * Token gather / scatter of the sparse encoders (SURVEY.md section 8(f) row 3;
 * models/sparse/unimodal_sparse_deformable_transformer.py:425-431 torch.gather over an index
 * repeated along the channels, :443-448 one `output[i].scatter(...)` per batch item in a Python
 * loop).  Rows of `row_bytes` bytes (C * element size), any element type.
 *   x, base, out_scatter [B, S, row]   idx [B, K] int64 (device)   src, out_gather [B, K, row]
 *   nums [B] int64 (device) or NULL: only the first nums[b] indices of item b count
 *   msda_token_gather : out[b,k] = x[b, idx[b,k]]   (k >= nums[b] or an index outside [0,S): zeros)
 *   msda_token_scatter: out[b,s] = src[b,k] where idx[b,k] == s for a counted k (indices of one item
 *                       are distinct, as torch.topk's are), else base[b,s]; zero_hit_rows != 0 writes
 *                       zeros instead of src rows (the gradient w.r.t. base; src may be NULL).
 *                       `inverse`: caller's scratch of B*S int32 (filled here).  out may not alias base.
 */
int msda_token_gather(const void* x, const int64_t* idx, const int64_t* nums, void* out,
                      int64_t B, int64_t S, int64_t K, int64_t row_bytes, void* stream);
int msda_token_scatter(const void* base, const int64_t* idx, const int64_t* nums, const void* src,
                       void* out, int* inverse, int64_t B, int64_t S, int64_t K, int64_t row_bytes,
                       int zero_hit_rows, void* stream);

/* Last CUDA runtime error code seen by the calling thread inside this library
 * (cudaError_t as int) and its string. */
int msda_last_cuda_error(void);
const char* msda_last_cuda_error_string(void);

#ifdef __cplusplus
}
#endif
#endif /* MSDA_B200_H_ */
