// msda_tile.cuh -- CTA-tile MSDeformAttn kernels for temporal (H == 1) maps: the
// encoder path of round 2.
//
// What changed against the warp-level row windows of msda_win.cuh:
//   * queries are processed in TIME ORDER, not in index order.  A small pre-pass
//     (msda_sort_kernel) bins the queries of each batch item by the mean x of their
//     level-0 points and writes a permutation.  The encoder's query list is the
//     concatenation of its levels (models/deformable/unimodal_deformable_transformer.py:
//     90-134), the sparse encoders hand the op score-ordered top-k tokens
//     (models/sparse/unimodal_sparse_deformable_transformer.py:218, 425-431): in time
//     order four neighbouring queries share their rows whatever the index order was
//     (18.8 packed rows per 4-query block at the encoder shape instead of 30.7).
//   * a TILE = 64 time-adjacent queries of one (batch, head), owned by a CTA of 128
//     threads.  The union of the rows the tile taps is, per level, one contiguous run
//     of value rows (~85 rows of 256 B at the encoder shape): it is brought to shared
//     memory ONCE by TMA (cp.async.bulk.tensor.2d boxes of 8 rows x D channels behind
//     an mbarrier, issued by one thread while the others build their plans) and all 16
//     blocks of the CTA read their rows with LDS.128.  No register ring of in-flight
//     global rows, no L1 dependence, value rows cross L2 -> SM ~1.3 times instead of ~6.
//   * the 4-query BLOCK of msda_win.cuh survives: 8 lanes, lane j owns channels
//     [4j, 4j+4) and [32+4j, 32+4j+4), weights merged per (row, query) in a small
//     shared-memory matrix, four queries sharing each row read.
// Tiles whose union window does not fit (random locations) and launches with H != 1
// run the per-tap body of msda_fast.cuh on the tile's query list.
//
// Semantics note (as msda_win.cuh): a row inside a block's window meets zero weights of
// the block's other queries, so a non-finite value row poisons up to 3 neighbours in time
// through 0 * inf; MSDA_FLAG_NO_WINDOW selects the per-tap kernels, which never read a row
// they do not tap.
#pragma once
#include <cooperative_groups.h>
#include <cuda.h>   // CUtensorMap (the type only; the encoder is resolved at run time)
#include <type_traits>

#include "msda_win.cuh"

namespace msda {

// tuning knobs (overridable at build time for A/B runs: build.sh -DMSDA_TILE_WARPS=8 ...)
#ifndef MSDA_TILE_WARPS
#define MSDA_TILE_WARPS 4
#endif
#ifndef MSDA_TILE_CAP
#define MSDA_TILE_CAP 128           // value rows per tile.  152 -> 128 (32 KB at fp32): 4 CTAs leave ~35 KB of L1
                                   // for the global loads; forward 55.8 -> 53.8 us, backward (4 CTAs) 96.0 -> 90.0
#endif
#ifndef MSDA_TILE_BLK
#define MSDA_TILE_BLK 56
#endif
#ifndef MSDA_TILE_WT_INTERLEAVE
#define MSDA_TILE_WT_INTERLEAVE 1   // 1: the 4 blocks of a warp interleave their weight rows (below)
#endif
#ifndef MSDA_TILE_BULK_RED
#define MSDA_TILE_BULK_RED 0        // 1: backward hands grad_value rows to the TMA unit (cp.reduce.async.bulk);
                                   // measured SLOWER (127 vs 96 us at ENC, profiles/r2c_sass_bulk_red.txt): off
#endif
#ifndef MSDA_TILE_RED_GROUP
#define MSDA_TILE_RED_GROUP 1       // 1: backward issues the reds of a row (all units) under ONE branch
#endif
#ifndef MSDA_TILE_SAVED_PLANS
#define MSDA_TILE_SAVED_PLANS 0     // 1: MSDA_FLAG_SAVE_PLAN is honoured (forward saves the tiles' plans, backward
                                   // starts from them).  Same results, no gain measured, and the extra branch
                                   // in the backward kernel slows its per-tap body by 38 %: off
#endif
#ifndef MSDA_TILE_ISSUE_SPLIT
#define MSDA_TILE_ISSUE_SPLIT 0     // 1: the TMA boxes of a tile are requested by 4 lanes (one per warp and level)
                                   // instead of thread 0.  Measured: no change at fp32, bf16 backward 83 -> 85 us
#endif
#ifndef MSDA_TILE_LEVEL_LOOP
#define MSDA_TILE_LEVEL_LOOP 0      // 1: the level loop around the row loops is a real loop (one copy of the body
                                   // instead of four).  Measured: forward unchanged, backward 86.1 -> 88.1 us
#endif
#ifndef MSDA_TILE_FWD_UNROLL
#define MSDA_TILE_FWD_UNROLL 2
#endif
#ifndef MSDA_TILE_FWD_CTAS
#define MSDA_TILE_FWD_CTAS 4
#endif
#ifndef MSDA_TILE_FWD_CTAS_BF16
#define MSDA_TILE_FWD_CTAS_BF16 6
#endif
#ifndef MSDA_TILE_BWD_CTAS
#define MSDA_TILE_BWD_CTAS 4
#endif
#ifndef MSDA_TILE_BWD_CTAS_BF16
#define MSDA_TILE_BWD_CTAS_BF16 5
#endif
constexpr int kTileWarps = MSDA_TILE_WARPS;
constexpr int kTileThreads = kTileWarps * 32;
constexpr int kTileBlocks = kTileThreads / 8;        // 4-query blocks per tile
constexpr int kTileQ = kTileBlocks * kWinQ;          // queries per tile (64 with 4 warps)
#ifndef MSDA_TILE_BOX
#define MSDA_TILE_BOX 8
#endif
constexpr int kTileBoxRows = MSDA_TILE_BOX;          // value rows per TMA box
constexpr int kTileFwdUnroll = MSDA_TILE_FWD_UNROLL; // unroll factor of the forward's row loops
constexpr int kTileCapRows = MSDA_TILE_CAP;          // value rows per tile (multiple of the box)
constexpr int kTileBlkRows = MSDA_TILE_BLK;          // packed rows per block (plan capacity)
// Layout of the merged weight matrices in shared memory.  Row t of block g (packed plan row; the
// packing is warp-uniform) is one float4 (one weight per query of the block):
//   interleaved: warp region [t][g & 3][qi] -- the 4 blocks of a warp sit in 4 different bank
//                groups, so the scatter's read-modify-writes and the backward's dot stores of a
//                warp conflict at most 2-way (levels h / h + 1 of one block);
//   private:     [g][t][qi], blocks 896 B apart = the same banks (4-5-way conflicts measured).
constexpr int kTileWtStride = MSDA_TILE_WT_INTERLEAVE ? 4 : 1;   // float4 units between rows t, t + 1
__device__ __forceinline__ int tile_wt_base(int g) {             // float4 index of (block g, row 0)
  return MSDA_TILE_WT_INTERLEAVE ? (g >> 2) * 4 * kTileBlkRows + (g & 3) : g * kTileBlkRows;
}
// A SAVED plan (MSDA_FLAG_SAVE_PLAN): what the forward's prologue worked out for a tile, kept in
// the workspace behind the query order so that the backward starts from it instead of the points.
// Per tile (index = blockIdx.x) kSaveStride bytes:
//   int [0..3]  first slab row of the tile's window per level   (s_win lo)
//   int [4..7]  last slab row                                    (s_win hi)
//   int [8..11] packed plan rows of warp 0..3 (rt; 64 B each in the weight region)
//   int [12]    1: the tile has a plan, 0: it runs the per-tap body (nothing else is valid)
//   +128 + 32 g: block g: int blo[4], int brows[4]
//   +kSaveHdrBytes: the warps' weight regions as they sit in shared memory (interleaved layout),
//                   rt * 64 bytes used of each
constexpr int kSaveHdrBytes = 128 + kTileBlocks * 32;
constexpr int kSaveWarpBytes = 4 * kTileBlkRows * 16;
constexpr int kSaveStride = kSaveHdrBytes + kTileBlocks * kTileBlkRows * 16;
static_assert(kSaveHdrBytes % 128 == 0 && kSaveStride % 128 == 0, "saved plans stay 128-byte aligned");
static_assert(kTileWarps == 4 || !MSDA_TILE_SAVED_PLANS, "a saved plan's header holds 4 warps");
constexpr int kSortBins = 4096;
constexpr int kSortThreads = 1024;
constexpr int kTileBig = 0x3fffffff;

// ---- optional phase timing (debug builds: -DMSDA_TILE_TIMING) -------------------
#ifdef MSDA_TILE_TIMING
__device__ unsigned long long g_tile_times[8 * 8192];
__device__ __forceinline__ void tile_stamp(int k) {
  if (threadIdx.x == 0 && blockIdx.x < 8192) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_tile_times[blockIdx.x * 8 + k] = t;
  }
}
__device__ unsigned long long g_sort_times[8 * 64];
__device__ __forceinline__ void sort_stamp(int k) {
  if (threadIdx.x == 0 && blockIdx.x < 64) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_sort_times[blockIdx.x * 8 + k] = t;
  }
}
#else
__device__ __forceinline__ void tile_stamp(int) {}
__device__ __forceinline__ void sort_stamp(int) {}
#endif

// ---- query order ---------------------------------------------------------------
// sum of the x coordinates of the 4 level-0 points of one (query, head)
template <typename T> struct SortKey;
template <> struct SortKey<float> {
  static __device__ __forceinline__ float sum4(const float* p) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    return (a.x + a.z) + (b.x + b.z);
  }
};
template <> struct SortKey<__nv_bfloat16> {
  static __device__ __forceinline__ float sum4(const __nv_bfloat16* p) {
    const uint4 l = __ldg(reinterpret_cast<const uint4*>(p));
    return (__uint_as_float(l.x << 16) + __uint_as_float(l.y << 16)) +
           (__uint_as_float(l.z << 16) + __uint_as_float(l.w << 16));
  }
};

// Where the sort key comes from: the op's sampling locations, or -- fused module front-end -- the raw
// offsets and reference points (x = ref + off * sc, sc = 1 / T_0 or ref1 * 0.5 / P).
template <typename T>
struct SortSrc {
  const T* loc;
  const T* off;
  const T* ref;
  const int64_t* shapes;
  int ref_dim;
};
template <typename T>
__device__ __forceinline__ float sort_key_sum4(const SortSrc<T>& s, const Dims& d, int b, int q, int hd) {
  if (s.off == nullptr)
    return SortKey<T>::sum4(s.loc + (((size_t)b * d.Lq + q) * d.M + hd) * (size_t)(d.L * d.P * 2));
  const T* op = s.off + (((size_t)b * d.Lq + q) * d.M + hd) * (size_t)(d.L * d.P);
  const T* rp = s.ref + ((size_t)b * d.Lq + q) * (size_t)(d.L * s.ref_dim);
  float o = 0.f;
#pragma unroll
  for (int p = 0; p < 4; ++p) o += (float)to_acc<T>(op[p]);
  const float sc = s.ref_dim == 1 ? 1.f / (float)s.shapes[1] : (float)to_acc<T>(rp[1]) * (0.5f / 4.f);
  return 4.f * (float)to_acc<T>(rp[0]) + o * sc;
}

// One CTA per batch item: counting sort of its Lq queries by time bin.  perm[b][i] = index of
// the i-th query in time order (ties in arbitrary order: no result depends on the order inside
// a bin, the forward is bitwise independent of the permutation).  P == 4.
// HK lanes share a query (one head each, the first HK = min(M, 2) heads), so all loads of a pass are in flight together; the bin of a query is kept in shared
// memory between the two passes when Lq allows it.
constexpr int kSortStash = 4096;   // static shared memory stays below 48 KB

template <typename T>
__global__ void __launch_bounds__(kSortThreads)
msda_sort_kernel(SortSrc<T> src, int* __restrict__ perm, Dims d, int hk_shift) {
  __shared__ int s_hist[kSortBins];
  __shared__ int s_warp[32];
  __shared__ unsigned s_bin[kSortStash];   // (bin << 16) | rank inside the bin
  // let the dependent (tile) kernel's CTAs be scheduled right away: they wait for this grid's
  // completion (griddepcontrol.wait) before they read the permutation
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int b = blockIdx.x, tid = threadIdx.x;
  sort_stamp(0);
  for (int i = tid; i < kSortBins; i += kSortThreads) s_hist[i] = 0;
  __syncthreads();
  sort_stamp(1);
  const int HK = 1 << hk_shift;
  const float scale = (float)kSortBins / (float)(4 * HK);
  const bool stash = d.Lq <= kSortStash;
  const int items = d.Lq << hk_shift;
  const int items_up = (items + kSortThreads - 1) / kSortThreads * kSortThreads;
  auto bin_of = [&](int item) -> int {   // all lanes of a warp call it together
    const int q = item >> hk_shift, hd = item & (HK - 1);
    float s = item < items ? sort_key_sum4<T>(src, d, b, q, hd) : 0.f;
    for (int o = 1; o < HK; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    // NaN -> bin 0 (fmaxf drops it), +-inf clamp
    return __float2int_rd(fminf(fmaxf(s * scale, 0.f), (float)(kSortBins - 1)));
  };
  // 4 items per thread and round: all loads of a round are issued before the first is used
  for (int base = tid; base < items_up; base += 4 * kSortThreads) {
    float ks[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int item = base + u * kSortThreads;
      ks[u] = item < items ? sort_key_sum4<T>(src, d, b, item >> hk_shift, item & (HK - 1)) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int item = base + u * kSortThreads;
      if (item - tid >= items_up) break;               // warp-uniform: the shuffles stay converged
      float sk = ks[u];
      for (int o = 1; o < HK; o <<= 1) sk += __shfl_xor_sync(0xffffffffu, sk, o);
      const int bin = __float2int_rd(fminf(fmaxf(sk * scale, 0.f), (float)(kSortBins - 1)));
      if (item < items && (item & (HK - 1)) == 0) {
        const int slot = atomicAdd(&s_hist[bin], 1);
        if (stash) s_bin[item >> hk_shift] = ((unsigned)bin << 16) | (unsigned)slot;
      }
    }
  }
  __syncthreads();
  sort_stamp(2);
  // exclusive scan of the histogram: 4 bins per thread, shuffle scan per warp, then over warps
  constexpr int PER = kSortBins / kSortThreads;
  int v[PER], sum = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k) { v[k] = s_hist[tid * PER + k]; sum += v[k]; }
  int inc = sum;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, s);
    if ((tid & 31) >= s) inc += t;
  }
  if ((tid & 31) == 31) s_warp[tid >> 5] = inc;
  __syncthreads();
  if (tid < 32) {
    int w = s_warp[tid], wi = w;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, wi, s);
      if (tid >= s) wi += t;
    }
    s_warp[tid] = wi - w;
  }
  __syncthreads();
  int run = s_warp[tid >> 5] + inc - sum;
#pragma unroll
  for (int k = 0; k < PER; ++k) { s_hist[tid * PER + k] = run; run += v[k]; }
  __syncthreads();
  sort_stamp(3);
  int* pb = perm + (size_t)b * d.Lq;
  if (stash) {   // position = start of the bin + the rank the first pass drew: no second atomic pass
    for (int q = tid; q < d.Lq; q += kSortThreads) pb[s_hist[s_bin[q] >> 16] + (int)(s_bin[q] & 0xffffu)] = q;
  } else {
#pragma unroll 4
    for (int item = tid; item < items_up; item += kSortThreads) {
      const int bin = bin_of(item);
      if (item < items && (item & (HK - 1)) == 0) pb[atomicAdd(&s_hist[bin], 1)] = item >> hk_shift;
    }
  }
  __syncthreads();
  sort_stamp(4);
}

// The same for Lq <= kSortStash, faster and DETERMINISTIC (ties in index order): a cluster of
// kSortCluster CTAs per batch item.  All of them read keys (the loads are one 32-byte sector per
// (query, head), fully divergent: one SM's L1 serves ~1 sector per clock, eight SMs eight), and
// drop w = (key16 << 16 | query) into the shared memory of the cluster's first CTA through DSMEM.
// That CTA counting-sorts the words on their top kSortBits bits and then ranks every word inside
// its bin (the words are unique), so the result does not depend on the order the atomics ran in.
constexpr int kSortCluster = 8;
constexpr int kSortBits = 11;
constexpr int kSortBins2 = 1 << kSortBits;

template <typename T>
__global__ void __cluster_dims__(kSortCluster, 1, 1) __launch_bounds__(kSortThreads)
msda_sort_cluster_kernel(SortSrc<T> src, int* __restrict__ perm, Dims d, int hk_shift) {
  namespace cg = cooperative_groups;
  __shared__ unsigned s_key[kSortStash];     // w per query, later the sorted words
  __shared__ unsigned s_tmp[kSortStash];     // (slot << 16 | bin) per query, later bin-ordered words
  __shared__ int s_hist[kSortBins2 + 1];
  __shared__ int s_warp[32];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int b = blockIdx.x / kSortCluster, tid = threadIdx.x;
  sort_stamp(0);
  const int HK = 1 << hk_shift;
  const float scale = 65536.f / (float)(4 * HK);
  unsigned* key0 = cluster.map_shared_rank(s_key, 0);
  const int per = (d.Lq + kSortCluster - 1) / kSortCluster;
  const int q0 = rank * per, nq = max(min(d.Lq - q0, per), 0);
  const int items = nq << hk_shift;
  const int items_up = (items + 31) & ~31;
  for (int item = tid; item < items_up; item += kSortThreads) {
    const int q = q0 + (item >> hk_shift), hd = item & (HK - 1);
    float sk = item < items ? sort_key_sum4<T>(src, d, b, q, hd) : 0.f;
    for (int o = 1; o < HK; o <<= 1) sk += __shfl_xor_sync(0xffffffffu, sk, o);
    // NaN -> 0 (fmaxf drops it), +-inf clamp
    const unsigned kb = (unsigned)__float2int_rd(fminf(fmaxf(sk * scale, 0.f), 65535.f));
    if (item < items && hd == 0) key0[q] = (kb << 16) | (unsigned)q;
  }
  if (rank == 0)
    for (int i = tid; i <= kSortBins2; i += kSortThreads) s_hist[i] = 0;
  cluster.sync();
  sort_stamp(1);
  if (rank != 0) return;
  // counts (order-independent) and a provisional slot per query
  for (int q = tid; q < d.Lq; q += kSortThreads) {
    const int bin = (int)(s_key[q] >> (32 - kSortBits));
    s_tmp[q] = ((unsigned)atomicAdd(&s_hist[bin], 1) << 16) | (unsigned)bin;
  }
  __syncthreads();
  // exclusive scan of the counts: kSortBins2 / kSortThreads bins per thread
  constexpr int PER = kSortBins2 / kSortThreads;
  int v[PER], sum = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k) { v[k] = s_hist[tid * PER + k]; sum += v[k]; }
  int inc = sum;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, s);
    if ((tid & 31) >= s) inc += t;
  }
  if ((tid & 31) == 31) s_warp[tid >> 5] = inc;
  __syncthreads();
  if (tid < 32) {
    int w = s_warp[tid], wi = w;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, wi, s);
      if (tid >= s) wi += t;
    }
    s_warp[tid] = wi - w;
  }
  __syncthreads();
  int run = s_warp[tid >> 5] + inc - sum;
#pragma unroll
  for (int k = 0; k < PER; ++k) { s_hist[tid * PER + k] = run; run += v[k]; }
  if (tid == kSortThreads - 1) s_hist[kSortBins2] = run;
  sort_stamp(2);
  // words into bin order (any order inside a bin), registers in between: s_tmp changes meaning
  unsigned wq[kSortStash / kSortThreads], pq[kSortStash / kSortThreads];
#pragma unroll
  for (int u = 0; u < kSortStash / kSortThreads; ++u) {
    const int q = tid + u * kSortThreads;
    if (q < d.Lq) { wq[u] = s_key[q]; pq[u] = s_tmp[q]; }
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < kSortStash / kSortThreads; ++u) {
    const int q = tid + u * kSortThreads;
    if (q < d.Lq) s_tmp[s_hist[pq[u] & 0xffffu] + (int)(pq[u] >> 16)] = wq[u];
  }
  __syncthreads();
  // rank inside the bin: the words are unique, so this fixes the order
  for (int p = tid; p < d.Lq; p += kSortThreads) {
    const unsigned w = s_tmp[p];
    const int bin = (int)(w >> (32 - kSortBits));
    const int lo = s_hist[bin], hi = s_hist[bin + 1];
    int r = 0;
    for (int x = lo; x < hi; ++x) r += s_tmp[x] < w;
    s_key[lo + r] = w;
  }
  __syncthreads();
  sort_stamp(3);
  int* pb = perm + (size_t)b * d.Lq;
  for (int i = tid; i < d.Lq; i += kSortThreads) pb[i] = (int)(s_key[i] & 0xffffu);
  sort_stamp(4);
}

// Loads of data the PRECEDING kernel of a programmatic dependent launch produced (the query order,
// the plan): plain ld.global behind griddepcontrol.wait.  A __ldg / const __restrict__ load is
// "invariant" to the compiler and may be hoisted above the wait.
__device__ __forceinline__ void tile_grid_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ int tile_ld_dep(const int* p) {
  int v;
  asm volatile("ld.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int4 tile_ld_dep4(const void* p) {
  int4 v;
  asm volatile("ld.global.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// ---- mbarrier / TMA ------------------------------------------------------------
__device__ __forceinline__ unsigned tile_smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void tile_mbar_init(unsigned long long* mbar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tile_smem_u32(mbar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tile_mbar_expect_tx(unsigned long long* mbar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tile_smem_u32(mbar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tile_mbar_wait(unsigned long long* mbar, unsigned parity) {
  const unsigned a = tile_smem_u32(mbar);
  unsigned ok, spins = 0;
  do {
    // a TMA box that never lands (a rejected descriptor) must fault, not hang the device
    if (++spins > (1u << 22)) __trap();
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(a), "r"(parity)
        : "memory");
  } while (!ok);
}
// one box (kTileBoxRows rows x D channels) of the value tensor viewed as [N*S rows][M*D]
__device__ __forceinline__ void tile_tma_load(void* smem_dst, const CUtensorMap* map, int col, int row,
                                              unsigned long long* mbar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(tile_smem_u32(smem_dst)), "l"(reinterpret_cast<unsigned long long>(map)), "r"(col), "r"(row),
      "r"(tile_smem_u32(mbar))
      : "memory");
}

// a unit of a value row from the shared-memory tile
// grad_value rows through the TMA unit: `bytes` of fp32 at `ssrc` (shared memory, written by generic
// stores + fence.proxy.async) are ADDED to global memory at `gdst` (both 16-byte aligned); the
// operation belongs to the issuing thread's current bulk group.  SASS: UBLKRED.
__device__ __forceinline__ void tile_bulk_red_f32(float* gdst, const void* ssrc, unsigned bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gdst),
               "r"(tile_smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
// bulk copy global -> shared (1-D), completion on an mbarrier
__device__ __forceinline__ void tile_bulk_load(void* smem_dst, const void* gsrc, unsigned bytes,
                                               unsigned long long* mbar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          tile_smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(tile_smem_u32(mbar))
      : "memory");
}
// bulk copy shared -> global (1-D) in the issuing thread's current bulk group
__device__ __forceinline__ void tile_bulk_store(void* gdst, const void* ssrc, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(tile_smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void tile_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tile_bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tile_fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

template <typename U>
__device__ __forceinline__ void tile_lds(const char* p, typename U::Raw (&buf)[U::RW]) {
  using Raw = typename U::Raw;
  if constexpr (U::RW == 4) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    if constexpr (std::is_same<Raw, float>::value) {
      buf[0] = __uint_as_float(t.x); buf[1] = __uint_as_float(t.y);
      buf[2] = __uint_as_float(t.z); buf[3] = __uint_as_float(t.w);
    } else {
      buf[0] = t.x; buf[1] = t.y; buf[2] = t.z; buf[3] = t.w;
    }
  } else {
    const uint2 t = *reinterpret_cast<const uint2*>(p);
    buf[0] = t.x; buf[1] = t.y;
  }
}

// ---- fused module front-end (second step) ----------
// With a TileFront the tile kernels take what MSDeformAttn.forward has BEFORE
// models/ops/modules/ms_deform_attn.py:98-117 -- raw sampling offsets, attention logits and
// reference points -- and do the softmax over L*P, x = ref + off / T_l (or the (centre, length)
// form) and y = 0.5 in their prologue: sampling_locations and attention_weights never exist in
// HBM.  The backward's epilogue turns (d/dx, d/d aw) straight into grad_offsets, grad_logits
// (softmax backward) and grad_ref (fp32 atomics over heads and points).
template <typename T>
struct TileFront {
  const T* off;        // [N, Lq, M, L, P]   (nullptr: not fused, the kernels read loc / aw)
  const T* logits;     // [N, Lq, M, L*P]
  const T* ref;        // [N, Lq, L, ref_dim]
  T* g_off;            // backward outputs
  T* g_logits;
  float* g_ref;        // [N, Lq, L, ref_dim] fp32, zeroed by the host side (nullptr: not wanted)
  const unsigned char* mask;   // [N, S] bool, true = padding row of value (nullptr: none): the row counts
                               // as zeros and gets no gradient (value.masked_fill, ms_deform_attn.py:96-97)
  int ref_dim;
};

// Padding mask on the value tile: flag the tile's rows, zero the masked ones (after the boxes have
// landed).  All threads; ends with a barrier.
template <int ROW_BYTES>
__device__ __forceinline__ void tile_apply_mask(const unsigned char* mrow, int S, const int (&tlo)[kWinMaxL],
                                                const int (&tbase)[kWinMaxL], int total, char* vtile,
                                                unsigned char* s_rowmask) {
  for (int t = threadIdx.x; t < total; t += kTileThreads) {
    int l = 0;
#pragma unroll
    for (int k = 1; k < kWinMaxL; ++k)
      if (t >= tbase[k]) l = k;
    const int slab = tlo[l] + (t - tbase[l]);
    const bool mk = slab >= 0 && slab < S && mrow[slab] != 0;
    s_rowmask[t] = mk ? 1 : 0;
    if (mk) {
      uint4* r = reinterpret_cast<uint4*>(vtile + (size_t)t * ROW_BYTES);
#pragma unroll
      for (int k = 0; k < ROW_BYTES / 16; ++k) r[k] = make_uint4(0, 0, 0, 0);
    }
  }
  __syncthreads();
}

template <typename T> struct Front4;       // 4 consecutive elements <-> 4 floats
template <> struct Front4<float> {
  static __device__ __forceinline__ void ld(const float* p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void st(float* p, const float (&v)[4]) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
  }
};
template <> struct Front4<__nv_bfloat16> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
    v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
    v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, const float (&v)[4]) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<const uint32_t*>(&a);
    t.y = *reinterpret_cast<const uint32_t*>(&b);
    __stcs(reinterpret_cast<uint2*>(p), t);
  }
};

// The 2 x 4 points of lane (qi, h) -- levels h and h + 2 of its query -- from the raw module
// tensors.  All 32 lanes call it together (the softmax sum crosses to lane ^ 4, which holds the
// query's other two levels).  off[][] are returned for the (centre, length) backward; sc[] is
// d x / d off per level.
template <typename T>
__device__ __forceinline__ void tile_front_points(const TileFront<T>& f, const Dims& d, size_t grp, size_t bq,
                                                  int h, const int* s_lvl, float (&x)[2][4], float (&y)[2][4],
                                                  float (&a)[2][4], float (&off)[2][4], float (&sc)[2]) {
  constexpr unsigned FULL = 0xffffffffu;
  float lg[2][4], r0[2];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int l = h + 2 * i;
    const bool on = l < d.L;
    const int ls = on ? l : 0;
    Front4<T>::ld(f.off + (grp * d.L + ls) * kWinP, off[i]);
    Front4<T>::ld(f.logits + (grp * d.L + ls) * kWinP, lg[i]);
    const T* rp = f.ref + (bq * d.L + ls) * f.ref_dim;
    r0[i] = (float)to_acc<T>(rp[0]);
    const float tl = (float)s_lvl[3 * ls + 1];
    sc[i] = f.ref_dim == 1 ? 1.f / tl : (float)to_acc<T>(rp[f.ref_dim - 1]) * (0.5f / (float)kWinP);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      if (!on) lg[i][p] = -INFINITY;
      mx = fmaxf(mx, lg[i][p]);
      // the module's own arithmetic (ms_deform_attn.py:102-108): ref + off / T_l, or
      // ref0 + off / P * ref1 * 0.5
      x[i][p] = f.ref_dim == 1 ? r0[i] + off[i][p] / tl
                               : r0[i] + off[i][p] / (float)kWinP * (float)to_acc<T>(rp[f.ref_dim - 1]) * 0.5f;
      y[i][p] = 0.5f;
    }
  }
  mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 4));
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      a[i][p] = expf(lg[i][p] - mx);     // exp(-inf) = 0 for the levels that do not exist
      sum += a[i][p];
    }
  sum += __shfl_xor_sync(FULL, sum, 4);
  const float inv = 1.f / sum;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int p = 0; p < 4; ++p) a[i][p] *= inv;
}

// ---- the tile's plan ------------------------------------------------------------
// Everything a thread knows after the prologue.  Block = the 8 lanes that share g = tid / 8;
// lane (qi = j & 3, h = j >> 2) owns the points of query qi at levels h and h + 2.
struct TilePlan {
  int blo[kWinMaxL];     // first slab row of the block's window per level
  int brows[kWinMaxL];   // rows of the block's window per level (0: none)
  int wrows[kWinMaxL];   // the largest brows of the warp's 4 blocks (PAIR: rounded up to even): the
                         // warp-uniform trip count; rows past brows carry zero weights
  int wmax[kWinMaxL];    // the same before rounding (what a forward over a PAIR plan loops to)
  int bpoff[kWinMaxL];   // packed position of the level inside the block's plan (warp-uniform)
  int trow[kWinMaxL];    // row of the shared-memory tile that holds slab row blo[l] (0 if none)
  int tlo[kWinMaxL];     // first slab row of the TILE's window per level
  int tbase[kWinMaxL];   // row of the shared-memory tile where that level starts
  int rt;                // packed rows of the block's plan (warp-uniform)
  int total;             // rows of the shared-memory tile in use (multiple of the box)
  bool ok;               // false: the tile runs the per-tap body
};

// a[l] for a run-time l without making `a` addressable (it stays in registers): the level loops of the
// row phases are NOT unrolled -- one copy of the loop body instead of four (code size, section 6 5p)
__device__ __forceinline__ int tile_sel4(int l, const int (&a)[kWinMaxL]) {
  static_assert(kWinMaxL == 4, "4 levels");
  return l == 0 ? a[0] : l == 1 ? a[1] : l == 2 ? a[2] : a[3];
}

// One thread: the value boxes of a tile (level l: rows tlo[l].. in boxes of kTileBoxRows, landing at
// tile row tbase[l]) -> TMA loads completing on mbar (whose expect_tx the caller has set)
// (levels l with l % nsplit == part: the warps of a CTA can share the issue, one elected lane each)
__device__ __forceinline__ void tile_issue_boxes(const CUtensorMap* vmap, const Dims& d, int b, int m,
                                                 const int (&tlo)[kWinMaxL], const int (&tbase)[kWinMaxL],
                                                 int total, char* vtile, unsigned row_bytes,
                                                 unsigned long long* mbar, int part = 0, int nsplit = 1) {
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    if (l % nsplit != part) continue;
    const int nb = ((l + 1 < kWinMaxL ? tbase[l + 1] : total) - tbase[l]) / kTileBoxRows;
#pragma unroll 1   // one thread runs this: keep it out of the kernel's code size
    for (int k = 0; k < nb; ++k)
      tile_tma_load(vtile + (size_t)(tbase[l] + k * kTileBoxRows) * row_bytes, vmap, m * d.D,
                    b * d.S + tlo[l] + k * kTileBoxRows, mbar);
  }
}

// Geometry of the thread's points, block and tile windows, TMA issue, weight scatter.
// PAIR: row loops run over pairs of rows (backward): every level of a plan is padded to an even
// row count.
// On return (plan.ok) the block's weights are in Wt, the value boxes are in flight on mbar.
// PRE: the thread's points were loaded by the caller (persistent kernel: a tile ahead), pre[i] = (x, y, a).
template <typename T, bool PAIR, bool ISSUE = true, bool FUSED = false, bool PRE = false>
__device__ __forceinline__ void tile_prologue(const CUtensorMap* vmap, const T* __restrict__ loc,
                                              const T* __restrict__ aw, const Dims& d, int b, int m,
                                              int q, const int* s_lvl, int* s_win, char* vtile,
                                              unsigned row_bytes, float4* Wt,
                                              unsigned long long* mbar, TilePlan& pl,
                                              const TileFront<T>* front = nullptr,
                                              const float (*pre)[3][4] = nullptr) {
  constexpr unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, j = tid & 7, qi = j & 3, h = j >> 2;
  const int gbase = (tid & 31) & ~7;
  const bool have_q = q >= 0;
  const int LP = d.L * kWinP;
  const size_t grp = ((size_t)b * d.Lq + (have_q ? q : 0)) * d.M + m;

  int r0[2][4];
  float w0[2][4], w1[2][4];
  int lo[2], hi[2];
  float fx[2][4], fy[2][4], fa[2][4];
  if (FUSED) {
    float foff[2][4], fsc[2];
    tile_front_points<T>(*front, d, grp, (size_t)b * d.Lq + (have_q ? q : 0), h, s_lvl, fx, fy, fa, foff, fsc);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int l = h + 2 * i;
    const bool active = have_q && l < d.L;
    const int ls = l < d.L ? l : 0;
    float x[4], y[4], a[4];
    if (FUSED) {
#pragma unroll
      for (int p = 0; p < 4; ++p) { x[p] = fx[i][p]; y[p] = fy[i][p]; a[p] = fa[i][p]; }
    } else if (PRE) {
#pragma unroll
      for (int p = 0; p < 4; ++p) { x[p] = pre[i][0][p]; y[p] = pre[i][1][p]; a[p] = pre[i][2][p]; }
    } else {
      WinPts<T>::ld(loc + (grp * LP + ls * kWinP) * 2, aw + grp * LP + ls * kWinP, x, y, a);
    }
    const int W = s_lvl[3 * ls + 1], start = s_lvl[3 * ls + 2];
    lo[i] = kTileBig; hi[i] = -kTileBig;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      WinCoef c;
      bool live, v0, v1;
      win_point<false>(x[p], y[p], a[p], W, start, d.border != 0, active, r0[i][p], w0[i][p], w1[i][p],
                       live, v0, v1, c);
      lo[i] = min(lo[i], live ? r0[i][p] : kTileBig);
      hi[i] = max(hi[i], live ? r0[i][p] + 1 : -kTileBig);
    }
    // a live point has -1 <= w_low <= W - 1: clip the pair to the map
    lo[i] = max(lo[i], start);
    hi[i] = min(hi[i], start + W - 1);
  }
  // the block's window per level: agree over its 4 queries
#pragma unroll
  for (int s = 1; s <= 2; s <<= 1)
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      lo[i] = min(lo[i], __shfl_xor_sync(FULL, lo[i], s));
      hi[i] = max(hi[i], __shfl_xor_sync(FULL, hi[i], s));
    }
  // the tile's window per level
  if (qi == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i)
      if (hi[i] >= lo[i]) {
        atomicMin(&s_win[h + 2 * i], lo[i]);
        atomicMax(&s_win[kWinMaxL + h + 2 * i], hi[i]);
      }
  }
  // every lane learns all levels of its block; pack them back to back
  int rt = 0;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int src = gbase + (l & 1) * 4;
    const int llo = __shfl_sync(FULL, lo[l >> 1], src);
    const int lhi = __shfl_sync(FULL, hi[l >> 1], src);
    const int rows = max(lhi - llo + 1, 0);
    const int wmax = __reduce_max_sync(FULL, rows);
    pl.blo[l] = llo;
    pl.brows[l] = rows;
    pl.wmax[l] = wmax;
    pl.wrows[l] = PAIR ? (wmax + 1) & ~1 : wmax;
    pl.bpoff[l] = rt;
    rt += pl.wrows[l];
  }
  tile_stamp(2);
  const bool over = __syncthreads_or(rt > kTileBlkRows) != 0;
  tile_stamp(3);

  int tlo[kWinMaxL], tbase[kWinMaxL];
  int total = 0;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int llo = s_win[l], lhi = s_win[kWinMaxL + l];
    const int rows = lhi >= llo ? lhi - llo + 1 : 0;
    tlo[l] = llo;
    tbase[l] = total;
    total += (rows + kTileBoxRows - 1) / kTileBoxRows * kTileBoxRows;
  }
  pl.total = total;
  pl.rt = rt;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) { pl.tlo[l] = tlo[l]; pl.tbase[l] = tbase[l]; }
  pl.ok = !over && total <= kTileCapRows;
  if (!pl.ok) return;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) pl.trow[l] = pl.brows[l] > 0 ? tbase[l] + pl.blo[l] - tlo[l] : 0;

  // thread 0 arrives with the byte count; the first lane of every warp requests the boxes of "its" levels
  // (a complete_tx that lands before the expect_tx only drives the count negative for a moment: the phase
  // cannot complete before thread 0's arrival)
#if MSDA_TILE_ISSUE_SPLIT
  if (ISSUE && total > 0) {
    if (tid == 0) tile_mbar_expect_tx(mbar, (unsigned)total * row_bytes);
    if ((tid & 31) == 0)
      tile_issue_boxes(vmap, d, b, m, tlo, tbase, total, vtile, row_bytes, mbar, tid >> 5, kTileWarps);
  }
#else
  if (ISSUE && tid == 0 && total > 0) {
    tile_mbar_expect_tx(mbar, (unsigned)total * row_bytes);
    tile_issue_boxes(vmap, d, b, m, tlo, tbase, total, vtile, row_bytes, mbar);
  }
#endif

  // the block's weight matrix: zero, then every lane scatters its own columns (level, qi)
  constexpr int WS = kTileWtStride;
  for (int t = j; t < rt; t += 8) Wt[t * WS] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncwarp();
  float* Wf = reinterpret_cast<float*>(Wt);
  int tb[2];
  tb[0] = h ? pl.bpoff[1] - pl.blo[1] : pl.bpoff[0] - pl.blo[0];
  tb[1] = h ? pl.bpoff[3] - pl.blo[3] : pl.bpoff[2] - pl.blo[2];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    float c0[2], c1[2];
    int t0[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      t0[i] = (r0[i][p] + tb[i]) * (4 * WS) + qi;
      c0[i] = w0[i][p] != 0.f ? Wf[t0[i]] : 0.f;
      c1[i] = w1[i][p] != 0.f ? Wf[t0[i] + 4 * WS] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      if (w0[i][p] != 0.f) Wf[t0[i]] = c0[i] + w0[i][p];
      if (w1[i][p] != 0.f) Wf[t0[i] + 4 * WS] = c1[i] + w1[i][p];
    }
  }
  __syncwarp();
}

// shared state every tile kernel sets up before the prologue; returns whether all maps have H == 1
__device__ __forceinline__ bool tile_setup(const int64_t* __restrict__ shapes,
                                           const int64_t* __restrict__ lsi,
                                           const int* __restrict__ perm, const Dims& d, int b, int chunk,
                                           int* s_lvl, int* s_win, int* s_q,
                                           unsigned long long* mbar) {
  const int tid = threadIdx.x;
  if (tid < 2 * kWinMaxL) s_win[tid] = tid < kWinMaxL ? kTileBig : -kTileBig;
  if (tid == 0) tile_mbar_init(mbar, 1);
  // programmatic dependent launch: this grid may start while the sort grid still runs; its
  // results (and everything before it in the stream) are visible after this wait
  tile_grid_wait();
  if (tid < kTileQ) {
    const int pos = chunk * kTileQ + tid;
    s_q[tid] = pos < d.Lq ? tile_ld_dep(perm + (size_t)b * d.Lq + pos) : -1;
  }
  return win_stage_levels(shapes, lsi, d.L, s_lvl);   // one barrier
}

// L2 prefetch of the points of the tile that will run ~2 waves later (tile index + kTileAhead): its
// CTA then finds its query list and its sampling points (or raw offsets / logits) in L2 instead of
// HBM -- the two dependent HBM round trips at the head of every tile CTA's life.
// Threads 64..127 of the CTA, after the grid-dependency wait (the permutation is final).
#ifndef MSDA_TILE_AHEAD
#define MSDA_TILE_AHEAD 0     // 1184 = two waves ahead.  Neutral with the 152-row tiles; with
                              // 128-row tiles and 4 CTAs both ways it costs: backward 90.0 vs 88.0 us.  Off
#endif
template <typename T>
__device__ __forceinline__ void tile_prefetch_ahead(const int* __restrict__ perm, const T* loc, const T* aw,
                                                    const TileFront<T>* front, const Dims& d, int nchunks) {
  if (MSDA_TILE_AHEAD <= 0) return;
  const int t = (int)threadIdx.x - 64;
  if (t < 0 || t >= kTileQ) return;
  const long long ft = (long long)blockIdx.x + MSDA_TILE_AHEAD;
  if (ft >= (long long)gridDim.x) return;
  const int fm = (int)(ft % d.M);
  const long long f2 = ft / d.M;
  const int fchunk = (int)(f2 % nchunks), fb = (int)(f2 / nchunks);
  const int pos = fchunk * kTileQ + t;
  if (pos >= d.Lq) return;
  const int q = perm[(size_t)fb * d.Lq + pos];
  const size_t grp = ((size_t)fb * d.Lq + q) * d.M + fm;
  const int LP = d.L * kWinP;
  const T* p0 = front && front->off ? front->off + grp * LP : loc + grp * LP * 2;
  const T* p1 = front && front->off ? front->logits + grp * LP : aw + grp * LP;
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p0));
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p1));
}

// The same for a SAVED plan: header and the first 2 KB of every warp's weight region of the tile
// that runs ~2 waves later (one 128-byte line per thread).
__device__ __forceinline__ void tile_prefetch_saved(const char* saved) {
  if (MSDA_TILE_AHEAD <= 0) return;
  const long long ft = (long long)blockIdx.x + MSDA_TILE_AHEAD;
  if (ft >= (long long)gridDim.x) return;
  const int t = threadIdx.x;
  const char* ts = saved + (size_t)ft * kSaveStride;
  const char* p;
  if (t < kSaveHdrBytes / 128) p = ts + t * 128;
  else if (t < kSaveHdrBytes / 128 + 4 * 16) {
    const int u = t - kSaveHdrBytes / 128;
    p = ts + kSaveHdrBytes + (u >> 4) * kSaveWarpBytes + (u & 15) * 128;
  } else return;
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// The per-tap bodies as out-of-line functions: they are the rare path and must not shape the
// register allocation of the tile path.
template <typename T, int NU, int NT>
__device__ __noinline__ void tile_fallback_fwd(const T* value, const T* loc, const T* aw, T* out, Dims d,
                                               int b, int m, const int* s_lvl, char* scratch,
                                               const int* s_q, const float* lpts = nullptr,
                                               const unsigned char* mrow = nullptr) {
  const int g = threadIdx.x >> 3;
  fwd_body<T, NU, 8, NT>(value, loc, aw, out, d, b, m, g, kTileQ, kTileBlocks, s_lvl,
                         reinterpret_cast<TapMeta<NT>*>(scratch) + g * 8, s_q, lpts, mrow);
}
template <typename T, int NU, int NT, typename GV, typename U>
__device__ __noinline__ void tile_fallback_bwd(const T* value, const T* loc, const T* aw, const T* gout,
                                               GV* gvalue, T* gloc, T* gaw, Dims d, int b, int m,
                                               const int* s_lvl, char* scratch, float gv_scale,
                                               const int* s_q, const float* lpts = nullptr,
                                               float* lgrad = nullptr, const unsigned char* mrow = nullptr) {
  const int g = threadIdx.x >> 3;
  bwd_body<T, NU, 8, NT, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, g, kTileQ, kTileBlocks,
                                s_lvl, reinterpret_cast<TapMeta<NT>*>(scratch) + g * 8, gv_scale, s_q, lpts,
                                lgrad, mrow);
}

// Fused front-end, tile without a plan: the thread's points (x, y, weight) as floats into shared
// memory [kTileQ][L*4][3], where the per-tap body finds them.  All threads; ends with a barrier.
template <typename T>
__device__ __forceinline__ void tile_front_stash(const TileFront<T>& f, const Dims& d, int b, int m, int q,
                                                 const int* s_lvl, float* lp) {
  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7, qi = j & 3, h = j >> 2;
  const int qq = q >= 0 ? q : 0;
  float x[2][4], y[2][4], a[2][4], off[2][4], sc[2];
  tile_front_points<T>(f, d, ((size_t)b * d.Lq + qq) * d.M + m, (size_t)b * d.Lq + qq, h, s_lvl, x, y, a, off, sc);
  const int LP = d.L * kWinP;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int l = h + 2 * i;
    if (l < d.L) {
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        float* o = lp + ((size_t)(g * kWinQ + qi) * LP + l * kWinP + p) * 3;
        o[0] = x[i][p]; o[1] = y[i][p]; o[2] = a[i][p];
      }
    }
  }
  __syncthreads();
}

// Fused front-end, backward epilogue: (d/dx, d/d weight) of the thread's 2 x 4 points ->
// grad_offsets, grad_logits (softmax backward over the query's L*4 points: this lane's 8 and lane
// ^ 4's), grad_ref (fp32 atomics: summed over points here, over heads in memory).  All 32 lanes.
template <typename T>
__device__ __forceinline__ void tile_front_backward(const TileFront<T>& f, const Dims& d, int b, int m, int q,
                                                    int h, const float (&a)[2][4], const float (&off)[2][4],
                                                    const float (&sc)[2], const float (&gx)[2][4],
                                                    const float (&ga)[2][4]) {
  float dot = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int p = 0; p < 4; ++p) dot = fmaf(a[i][p], ga[i][p], dot);
  dot += __shfl_xor_sync(0xffffffffu, dot, 4);
  if (q < 0) return;
  const size_t grp = ((size_t)b * d.Lq + q) * d.M + m, bq = (size_t)b * d.Lq + q;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int l = h + 2 * i;
    if (l >= d.L) continue;
    float go[4], gl[4], gr0 = 0.f, gr1 = 0.f;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      go[p] = gx[i][p] * sc[i];
      gl[p] = a[i][p] * (ga[i][p] - dot);
      gr0 += gx[i][p];
      gr1 = fmaf(gx[i][p], off[i][p], gr1);
    }
    Front4<T>::st(f.g_off + (grp * d.L + l) * kWinP, go);
    Front4<T>::st(f.g_logits + (grp * d.L + l) * kWinP, gl);
    if (f.g_ref) {
      atomicAdd(f.g_ref + (bq * d.L + l) * f.ref_dim, gr0);
      if (f.ref_dim == 2) atomicAdd(f.g_ref + (bq * d.L + l) * 2 + 1, gr1 * (0.5f / (float)kWinP));
    }
  }
}

// Forward, after the prologue: the tile's plan -> its slot `ts` of the workspace (layout at
// kSaveStride).  All threads; the warps' weight regions leave as bulk copies (the caller waits for
// their reads before the CTA exits).
__device__ __forceinline__ void tile_save_plan(char* ts, const TilePlan& pl, const int* s_win, char* s_plan) {
  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7, warp = tid >> 5;
  if (j == 0) {
    int4* hb = reinterpret_cast<int4*>(ts + 128 + g * 32);
    hb[0] = make_int4(pl.blo[0], pl.blo[1], pl.blo[2], pl.blo[3]);
    hb[1] = make_int4(pl.brows[0], pl.brows[1], pl.brows[2], pl.brows[3]);
  }
  const int rt = pl.ok ? pl.rt : 0;
  if ((tid & 31) == 0) reinterpret_cast<int*>(ts)[8 + warp] = rt;
  if (tid == 0) {
    int4* h = reinterpret_cast<int4*>(ts);
    h[0] = make_int4(s_win[0], s_win[1], s_win[2], s_win[3]);
    h[1] = make_int4(s_win[4], s_win[5], s_win[6], s_win[7]);
    reinterpret_cast<int*>(ts)[12] = pl.ok ? 1 : 0;
  }
  if (rt > 0) {
    tile_fence_async_smem();   // the scatter's generic-proxy writes -> visible to the bulk copy
    __syncwarp();
    if ((tid & 31) == 0) {
      tile_bulk_store(ts + kSaveHdrBytes + warp * kSaveWarpBytes, s_plan + warp * kSaveWarpBytes,
                      (unsigned)rt * 64u);
      tile_bulk_commit();
    }
  }
}

// Backward from a saved plan: the header -> TilePlan (as tile_prologue<PAIR = true> leaves it);
// thread 0 requests the value boxes and the warps' weight regions on mbar.
__device__ __forceinline__ void tile_load_plan(const char* ts, const CUtensorMap* vmap, const Dims& d, int b,
                                               int m, char* vtile, unsigned row_bytes, char* s_plan,
                                               unsigned long long* mbar, TilePlan& pl) {
  constexpr unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, g = tid >> 3;
  const int4* hb = reinterpret_cast<const int4*>(ts + 128 + g * 32);
  const int4 b0 = __ldg(hb), b1 = __ldg(hb + 1);
  const int4 lo4 = __ldg(reinterpret_cast<const int4*>(ts)), hi4 = __ldg(reinterpret_cast<const int4*>(ts) + 1);
  const int blo[kWinMaxL] = {b0.x, b0.y, b0.z, b0.w}, brows[kWinMaxL] = {b1.x, b1.y, b1.z, b1.w};
  const int lo[kWinMaxL] = {lo4.x, lo4.y, lo4.z, lo4.w}, hi[kWinMaxL] = {hi4.x, hi4.y, hi4.z, hi4.w};
  int rt = 0, total = 0;
  int tlo[kWinMaxL], tbase[kWinMaxL];
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int wmax = __reduce_max_sync(FULL, brows[l]);
    pl.blo[l] = blo[l];
    pl.brows[l] = brows[l];
    pl.wmax[l] = wmax;
    pl.wrows[l] = (wmax + 1) & ~1;
    pl.bpoff[l] = rt;
    rt += pl.wrows[l];
    const int rows = hi[l] >= lo[l] ? hi[l] - lo[l] + 1 : 0;
    tlo[l] = lo[l];
    tbase[l] = total;
    total += (rows + kTileBoxRows - 1) / kTileBoxRows * kTileBoxRows;
  }
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    pl.tlo[l] = tlo[l];
    pl.tbase[l] = tbase[l];
    pl.trow[l] = brows[l] > 0 ? tbase[l] + blo[l] - tlo[l] : 0;
  }
  pl.rt = rt;
  pl.total = total;
  pl.ok = true;
  if (tid == 0 && total > 0) {
    const int4 r4 = __ldg(reinterpret_cast<const int4*>(ts) + 2);
    const int rts[4] = {r4.x, r4.y, r4.z, r4.w};
    tile_mbar_expect_tx(mbar, (unsigned)total * row_bytes + (unsigned)(rts[0] + rts[1] + rts[2] + rts[3]) * 64u);
    tile_issue_boxes(vmap, d, b, m, tlo, tbase, total, vtile, row_bytes, mbar);
#pragma unroll
    for (int w = 0; w < 4; ++w)
      if (rts[w] > 0)
        tile_bulk_load(s_plan + w * kSaveWarpBytes, ts + kSaveHdrBytes + w * kSaveWarpBytes,
                       (unsigned)rts[w] * 64u, mbar);
  }
}

template <typename T, int NU>
struct TileSmem {
  static constexpr int kRowBytes = NU * 8 * 16;                        // D * sizeof(T)
  static constexpr int kValueBytes = kTileCapRows * kRowBytes;
  static constexpr int kPlanBytes = kTileBlocks * kTileBlkRows * 16;
#ifndef MSDA_TILE_SMEM_PAD
#define MSDA_TILE_SMEM_PAD 0      // experiment knob: extra bytes per CTA, to cap the CTAs per SM
#endif
  static constexpr int kBytes = kValueBytes + kPlanBytes + 128 + MSDA_TILE_SMEM_PAD;
};

// Backward: + the staging rows of the bulk reductions, [2 slots][block][2 rows][one head's row of
// grad_value], when grad_value accumulates as fp32 (not the fixed-point image of the deterministic form)
template <typename GV>
struct TileBulkRed { static constexpr bool value = MSDA_TILE_BULK_RED != 0 && std::is_same<GV, float>::value; };
template <typename T, int NU, typename GV, typename U>
constexpr size_t tile_bwd_smem() {
  return (size_t)kTileCapRows * NU * 8 * U::BYTES + (size_t)kTileBlocks * kTileBlkRows * 16 + 128 +
         (TileBulkRed<GV>::value ? (size_t)2 * kTileBlocks * 2 * NU * 8 * U::E * 4 : 0) + MSDA_TILE_SMEM_PAD;
}

// ------------------------------- forward ------------------------------------
// CTAs per SM the forward is compiled for: 4 at fp32 (48 KB of shared memory each), 6 for bf16 rows (half
// the tile bytes; 80 registers: ENC 49.6 -> 47.6 (5 CTAs) -> 43.6 us (6), MM_VV 39.4 -> 37.4 us)
template <typename T>
struct TileFwdCtas { static constexpr int value = std::is_same<T, __nv_bfloat16>::value ? MSDA_TILE_FWD_CTAS_BF16 : MSDA_TILE_FWD_CTAS; };

template <typename T, int NU, bool FUSED = false>
__global__ void __launch_bounds__(kTileThreads, TileFwdCtas<T>::value)
msda_fwd_tile_kernel(const __grid_constant__ CUtensorMap vmap, const T* __restrict__ value,
                     const int64_t* __restrict__ shapes, const int64_t* __restrict__ lsi,
                     const T* __restrict__ loc, const T* __restrict__ aw,
                     const int* __restrict__ perm, T* __restrict__ out, Dims d, int nchunks,
                     TileFront<T> front, char* save) {
  using U = Unit<T>;
  using Raw = typename U::Raw;
  using SM = TileSmem<T, NU>;
  constexpr int G = 8;
  constexpr int E = U::E;
  constexpr int RB = SM::kRowBytes;
  __shared__ int s_lvl[3 * kWinMaxL];
  __shared__ int s_win[2 * kWinMaxL];
  __shared__ int s_q[kTileQ];
  __shared__ __align__(8) unsigned long long s_mbar;
  extern __shared__ __align__(128) char s_raw[];   // TMA destination: 128-byte aligned
  char* vtile = s_raw;
  char* s_plan = vtile + SM::kValueBytes;

  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7;
  const int m = blockIdx.x % d.M;
  const int t2 = blockIdx.x / d.M;
  const int chunk = t2 % nchunks, b = t2 / nchunks;
  tile_stamp(0);
  const bool one_d = tile_setup(shapes, lsi, perm, d, b, chunk, s_lvl, s_win, s_q, &s_mbar);
  tile_stamp(1);
  tile_prefetch_ahead<T>(perm, loc, aw, FUSED ? &front : nullptr, d, nchunks);
  if (!one_d) {
    if (FUSED) __trap();   // the module's maps are (1, T_l): precondition of the fused entry points
    tile_fallback_fwd<T, NU, 4>(value, loc, aw, out, d, b, m, s_lvl, s_plan, s_q);
    return;
  }
  float4* Wt = reinterpret_cast<float4*>(s_plan) + tile_wt_base(g);
  TilePlan pl;
  // the plan is laid out as the backward wants it (levels padded to even row counts), so that a
  // saved plan serves both; this kernel's loops stop at wmax
  tile_prologue<T, true, true, FUSED>(&vmap, loc, aw, d, b, m, s_q[g * kWinQ + (j & 3)], s_lvl, s_win, vtile,
                                      RB, Wt, &s_mbar, pl, &front);
#if !MSDA_TILE_SAVED_PLANS
  save = nullptr;
#endif
  if (save) tile_save_plan(save + (size_t)blockIdx.x * kSaveStride, pl, s_win, s_plan);
  if (!pl.ok) {
    if (FUSED) {
      float* lp = reinterpret_cast<float*>(vtile);
      tile_front_stash<T>(front, d, b, m, s_q[g * kWinQ + (j & 3)], s_lvl, lp);
      tile_fallback_fwd<T, NU, 2>(value, loc, aw, out, d, b, m, s_lvl, s_plan + SM::kPlanBytes - 4096, s_q, lp,
                                  front.mask ? front.mask + (size_t)b * d.S : nullptr);
    } else {
      tile_fallback_fwd<T, NU, 2>(value, loc, aw, out, d, b, m, s_lvl, s_plan, s_q);
    }
    return;
  }
  tile_stamp(4);
  if (pl.total > 0) tile_mbar_wait(&s_mbar, 0);
  tile_stamp(5);
  if (FUSED && front.mask) {
    __shared__ unsigned char s_rowmask[kTileCapRows];
    tile_apply_mask<RB>(front.mask + (size_t)b * d.S, d.S, pl.tlo, pl.tbase, pl.total, vtile, s_rowmask);
  }

  float2 acc[kWinQ][NU][E / 2];
#pragma unroll
  for (int q = 0; q < kWinQ; ++q)
#pragma unroll
    for (int uu = 0; uu < NU; ++uu)
#pragma unroll
      for (int c = 0; c < E / 2; ++c) acc[q][uu][c] = make_float2(0.f, 0.f);

  // warp-uniform row loops: a block past its own rows re-reads its last row against zero weights.
  // (A hand-pipelined form -- weights and row of step i + 1 requested before the FFMA2 of step i --
  // measured 2 us slower at ENC: DESIGN.md section 6.)
#if MSDA_TILE_LEVEL_LOOP
#pragma unroll 1
#else
#pragma unroll
#endif
  for (int l = 0; l < kWinMaxL; ++l) {
    const char* vr = vtile + tile_sel4(l, pl.trow) * RB + j * 16;
    const float4* wl = Wt + tile_sel4(l, pl.bpoff) * kTileWtStride;
    const int n = tile_sel4(l, pl.wmax), nb = tile_sel4(l, pl.brows);
#pragma unroll kTileFwdUnroll
    for (int i = 0; i < n; ++i) {
      const float4 wv = wl[i * kTileWtStride];
      const float wq[kWinQ] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
      for (int uu = 0; uu < NU; ++uu) {
        Raw raw[U::RW];
        tile_lds<U>(vr + uu * (G * 16), raw);
        float v[E];
        U::unpack_raw(raw, v);
#pragma unroll
        for (int q = 0; q < kWinQ; ++q) {
          const float2 w2 = make_float2(wq[q], wq[q]);
#pragma unroll
          for (int c = 0; c < E / 2; ++c)
            acc[q][uu][c] = __ffma2_rn(w2, make_float2(v[2 * c], v[2 * c + 1]), acc[q][uu][c]);
        }
      }
      if (i + 1 < nb) vr += RB;
    }
  }
  tile_stamp(6);
#pragma unroll
  for (int q = 0; q < kWinQ; ++q) {
    const int qq = s_q[g * kWinQ + q];
    if (qq >= 0) {
      T* og = out + (((size_t)b * d.Lq + qq) * d.M + m) * d.D + j * E;
#pragma unroll
      for (int uu = 0; uu < NU; ++uu) {
        float o[E];
#pragma unroll
        for (int c = 0; c < E / 2; ++c) { o[2 * c] = acc[q][uu][c].x; o[2 * c + 1] = acc[q][uu][c].y; }
        U::st_stream(og + uu * (G * E), o);
      }
    }
  }
  tile_stamp(7);
  if (save && (tid & 31) == 0) tile_bulk_wait_read<0>();   // shared memory must outlive the bulk copy's reads
}

// ------------------------- forward, persistent form (experiment) --------------------------
// One CTA walks tiles blockIdx.x, + gridDim.x, ...  While it works on a tile it already holds the
// NEXT tile's query list (shared memory, double buffered) and, from the middle of the prologue on, the
// next tile's sampling points in registers -- the two dependent global round trips at the head of a
// tile (perm, then points) are off its critical path.  fp32, unfused front-end; same results as
// msda_fwd_tile_kernel (same prologue, same row loops).  MSDA_FLAG_PERSIST.
template <int NU>
__global__ void __launch_bounds__(kTileThreads, MSDA_TILE_FWD_CTAS)
msda_fwd_tile_persist_kernel(const __grid_constant__ CUtensorMap vmap, const float* __restrict__ value,
                             const int64_t* __restrict__ shapes, const int64_t* __restrict__ lsi,
                             const float* __restrict__ loc, const float* __restrict__ aw,
                             const int* __restrict__ perm, float* __restrict__ out, Dims d, int nchunks,
                             int ntiles) {
  using T = float;
  using U = Unit<T>;
  using Raw = typename U::Raw;
  using SM = TileSmem<T, NU>;
  constexpr int G = 8;
  constexpr int E = U::E;
  constexpr int RB = SM::kRowBytes;
  __shared__ int s_lvl[3 * kWinMaxL];
  __shared__ int s_win2[2][2 * kWinMaxL];
  __shared__ int s_q2[2][kTileQ];
  __shared__ __align__(8) unsigned long long s_mbar;
  extern __shared__ __align__(128) char s_raw[];
  char* vtile = s_raw;
  char* s_plan = vtile + SM::kValueBytes;

  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7, qi = j & 3, h = j >> 2;
  const int LP = d.L * kWinP;
  if (tid < 2 * kWinMaxL) {
    s_win2[0][tid] = tid < kWinMaxL ? kTileBig : -kTileBig;
    s_win2[1][tid] = tid < kWinMaxL ? kTileBig : -kTileBig;
  }
  if (tid == 0) tile_mbar_init(&s_mbar, 1);
  tile_grid_wait();
  auto tile_coords = [&](int tile, int& b, int& m, int& chunk) {
    m = tile % d.M;
    const int t2 = tile / d.M;
    chunk = t2 % nchunks;
    b = t2 / nchunks;
  };
  auto stage_queries = [&](int tile, int* sq) {   // threads 0..63: the tile's query list
    if (tid < kTileQ) {
      int b, m, chunk;
      tile_coords(tile, b, m, chunk);
      const int pos = chunk * kTileQ + tid;
      sq[tid] = pos < d.Lq ? tile_ld_dep(perm + (size_t)b * d.Lq + pos) : -1;
    }
  };
  // the thread's points of a tile: (x, y, a) of 4 points at levels h and h + 2
  auto load_points = [&](int tile, int q, float (&pt)[2][3][4]) {
    int b, m, chunk;
    tile_coords(tile, b, m, chunk);
    const size_t grp = ((size_t)b * d.Lq + (q >= 0 ? q : 0)) * d.M + m;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int l = h + 2 * i, ls = l < d.L ? l : 0;
      WinPts<T>::ld(loc + (grp * LP + ls * kWinP) * 2, aw + grp * LP + ls * kWinP, pt[i][0], pt[i][1], pt[i][2]);
    }
  };
  int tile = blockIdx.x;
  stage_queries(tile, s_q2[0]);
  const bool one_d = win_stage_levels(shapes, lsi, d.L, s_lvl);   // barrier: s_q2[0], s_win2, mbar visible
  if (!one_d) {
    for (int it = 0; tile < ntiles; tile += gridDim.x, ++it) {
      int b, m, chunk;
      tile_coords(tile, b, m, chunk);
      if (it > 0) { stage_queries(tile, s_q2[0]); __syncthreads(); }
      tile_fallback_fwd<T, NU, 4>(value, loc, aw, out, d, b, m, s_lvl, s_plan, s_q2[0]);
      __syncthreads();
    }
    return;
  }
  float pts[2][3][4];
  load_points(tile, s_q2[0][g * kWinQ + qi], pts);
  float4* Wt = reinterpret_cast<float4*>(s_plan) + tile_wt_base(g);
  unsigned phase = 0;
  for (int it = 0; tile < ntiles; ++it) {
    const int cur = it & 1, next = tile + (int)gridDim.x;
    int b, m, chunk;
    tile_coords(tile, b, m, chunk);
    int* s_q = s_q2[cur];
    int* s_win = s_win2[cur];
    if (next < ntiles) stage_queries(next, s_q2[cur ^ 1]);   // visible after the prologue's barrier
    TilePlan pl;
    tile_prologue<T, true, true, false, true>(&vmap, loc, aw, d, b, m, s_q[g * kWinQ + qi], s_lvl, s_win, vtile, RB,
                                              Wt, &s_mbar, pl, nullptr, pts);
    // the other window buffer is free again (its readers passed the previous tile's last barrier)
    if (tid < 2 * kWinMaxL) s_win2[cur ^ 1][tid] = tid < kWinMaxL ? kTileBig : -kTileBig;
    // next tile's points: requested now, used in the next iteration
    float nxt[2][3][4];
    if (next < ntiles) load_points(next, s_q2[cur ^ 1][g * kWinQ + qi], nxt);
    if (!pl.ok) {
      tile_fallback_fwd<T, NU, 2>(value, loc, aw, out, d, b, m, s_lvl, s_plan, s_q);
    } else {
      if (pl.total > 0) { tile_mbar_wait(&s_mbar, phase); phase ^= 1u; }
      float2 acc[kWinQ][NU][E / 2];
#pragma unroll
      for (int q = 0; q < kWinQ; ++q)
#pragma unroll
        for (int uu = 0; uu < NU; ++uu)
#pragma unroll
          for (int c = 0; c < E / 2; ++c) acc[q][uu][c] = make_float2(0.f, 0.f);
#pragma unroll
      for (int l = 0; l < kWinMaxL; ++l) {
        const char* vr = vtile + pl.trow[l] * RB + j * 16;
        const float4* wl = Wt + pl.bpoff[l] * kTileWtStride;
        const int n = pl.wmax[l], nb = pl.brows[l];
#pragma unroll kTileFwdUnroll
        for (int i = 0; i < n; ++i) {
          const float4 wv = wl[i * kTileWtStride];
          const float wq[kWinQ] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
          for (int uu = 0; uu < NU; ++uu) {
            Raw raw[U::RW];
            tile_lds<U>(vr + uu * (G * 16), raw);
            float v[E];
            U::unpack_raw(raw, v);
#pragma unroll
            for (int q = 0; q < kWinQ; ++q) {
              const float2 w2 = make_float2(wq[q], wq[q]);
#pragma unroll
              for (int c = 0; c < E / 2; ++c)
                acc[q][uu][c] = __ffma2_rn(w2, make_float2(v[2 * c], v[2 * c + 1]), acc[q][uu][c]);
            }
          }
          if (i + 1 < nb) vr += RB;
        }
      }
#pragma unroll
      for (int q = 0; q < kWinQ; ++q) {
        const int qq = s_q[g * kWinQ + q];
        if (qq >= 0) {
          T* og = out + (((size_t)b * d.Lq + qq) * d.M + m) * d.D + j * E;
#pragma unroll
          for (int uu = 0; uu < NU; ++uu) {
            float o[E];
#pragma unroll
            for (int c = 0; c < E / 2; ++c) { o[2 * c] = acc[q][uu][c].x; o[2 * c + 1] = acc[q][uu][c].y; }
            U::st_stream(og + uu * (G * E), o);
          }
        }
      }
    }
    __syncthreads();   // the tile, the weights and this query list may be overwritten
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int p = 0; p < 4; ++p) pts[i][k][p] = nxt[i][k][p];
    tile = next;
  }
}

// ------------------------------- backward -----------------------------------
// Rows come from the shared-memory tile; grad_value rows leave as 128-bit reds (one per row,
// unit and lane of a block, 4.7 rows per query in time order); the dots <grad_out, value row>
// are transpose-reduced over the 8 lanes for two rows at a time and stored over the weights;
// the epilogue reads the points again (L2) and forms grad_attn_weight / grad_sampling_loc.
// CTAs per SM the backward is compiled for (fp32: 4 = 128 registers; with 152-row value tiles it was
// faster at 3 -- no L1 left at 4; bf16, whose tile is half the bytes: 5 = 96 registers, ENC 89.1 -> 87.0 us)
template <typename T>
struct TileBwdCtas { static constexpr int value = std::is_same<T, __nv_bfloat16>::value ? MSDA_TILE_BWD_CTAS_BF16 : MSDA_TILE_BWD_CTAS; };

template <typename T, int NU, typename GV = float, typename U = Unit<T>, bool FUSED = false>
__global__ void __launch_bounds__(kTileThreads, TileBwdCtas<T>::value)
msda_bwd_tile_kernel(const __grid_constant__ CUtensorMap vmap, const T* __restrict__ value,
                     const int64_t* __restrict__ shapes, const int64_t* __restrict__ lsi,
                     const T* __restrict__ loc, const T* __restrict__ aw,
                     const T* __restrict__ gout, const int* __restrict__ perm,
                     GV* __restrict__ gvalue, T* __restrict__ gloc, T* __restrict__ gaw, Dims d,
                     int nchunks, const DetHeader* __restrict__ det, TileFront<T> front,
                     const char* __restrict__ saved) {
  const float gv_scale = det ? (float)det->scale : 1.f;
  using Raw = typename U::Raw;
  constexpr int G = 8;
  constexpr int E = U::E;
  constexpr int RB = NU * G * U::BYTES;          // D * sizeof(T)
  constexpr int kValueBytes = kTileCapRows * RB;
  __shared__ int s_lvl[3 * kWinMaxL];
  __shared__ int s_win[2 * kWinMaxL];
  __shared__ int s_q[kTileQ];
  __shared__ __align__(8) unsigned long long s_mbar;
  extern __shared__ __align__(128) char s_raw[];   // TMA destination: 128-byte aligned
  char* vtile = s_raw;
  char* s_plan = vtile + kValueBytes;

  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7, qi = j & 3, h = j >> 2;
  const int m = blockIdx.x % d.M;
  const int t2 = blockIdx.x / d.M;
  const int chunk = t2 % nchunks, b = t2 / nchunks;
  tile_stamp(0);
  const bool one_d = tile_setup(shapes, lsi, perm, d, b, chunk, s_lvl, s_win, s_q, &s_mbar);
  tile_stamp(1);
  tile_prefetch_ahead<T>(perm, loc, aw, FUSED ? &front : nullptr, d, nchunks);
  if (!one_d) {
    if (FUSED) __trap();
    tile_fallback_bwd<T, NU, 4, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, s_lvl, s_plan,
                                       gv_scale, s_q);
    return;
  }
  // grad_out of the block's 4 queries (this lane's channels): requested first, used after the
  // prologue, whose work covers their way from HBM
  Raw gor[kWinQ][NU][U::RW];
  {
    const unsigned qstride = (unsigned)(d.M * d.D);
    const T* gb = gout + ((size_t)b * d.Lq * d.M + m) * d.D + j * E;
#pragma unroll
    for (int q = 0; q < kWinQ; ++q) {
      const int qq = s_q[g * kWinQ + q];
#pragma unroll
      for (int uu = 0; uu < NU; ++uu) {
#pragma unroll
        for (int k = 0; k < U::RW; ++k) gor[q][uu][k] = Raw(0);
        U::ldg_keep_stream(reinterpret_cast<const char*>(gb + (size_t)(qq >= 0 ? qq : 0) * qstride +
                                                         uu * (G * E)),
                           qq >= 0, gor[q][uu]);
      }
    }
  }
  const int my_q = s_q[g * kWinQ + qi];
  float4* Wt = reinterpret_cast<float4*>(s_plan) + tile_wt_base(g);
  float* Gs = reinterpret_cast<float*>(Wt);   // the dots G[t][qi] overwrite Wt[t][qi] pair by pair
  TilePlan pl;
  // the forward's saved plan of this tile (same sampling points), if there is one and the tile has a
  // plan at all: windows and merged weights come from the workspace, not from the points
#if MSDA_TILE_SAVED_PLANS
  const char* ts = saved ? saved + (size_t)blockIdx.x * kSaveStride : nullptr;
#else
  const char* ts = nullptr;
#endif
  if (ts) tile_prefetch_saved(saved);
  if (ts && __ldg(reinterpret_cast<const int*>(ts) + 12) != 0) {
    tile_stamp(2);
    tile_load_plan(ts, &vmap, d, b, m, vtile, RB, s_plan, &s_mbar, pl);
    tile_stamp(3);
  } else
    tile_prologue<T, true, true, FUSED>(&vmap, loc, aw, d, b, m, my_q, s_lvl, s_win, vtile, RB, Wt, &s_mbar, pl,
                                        &front);
  if (!pl.ok) {
    if (FUSED) {
      // points to shared memory, per-tap body with its per-point gradients to shared memory, then
      // the fused epilogue on those
      const int LPf = d.L * kWinP;
      float* lp = reinterpret_cast<float*>(vtile);
      float* lg = lp + kTileQ * LPf * 3;
      tile_front_stash<T>(front, d, b, m, my_q, s_lvl, lp);
      tile_fallback_bwd<T, NU, 2, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, s_lvl,
                                         s_plan + kTileBlocks * kTileBlkRows * 16 - 4096, gv_scale, s_q, lp, lg,
                                         front.mask ? front.mask + (size_t)b * d.S : nullptr);
      __syncthreads();
      float x[2][4], y[2][4], a[2][4], off[2][4], sc[2], gxx[2][4], gaa[2][4];
      const int qs = my_q >= 0 ? my_q : 0;
      tile_front_points<T>(front, d, ((size_t)b * d.Lq + qs) * d.M + m, (size_t)b * d.Lq + qs, h, s_lvl, x, y, a,
                           off, sc);
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const int l = h + 2 * i;
          const float* gq = lg + ((size_t)(g * kWinQ + qi) * LPf + (l < d.L ? l : 0) * kWinP + p) * 3;
          gxx[i][p] = (l < d.L && my_q >= 0) ? gq[0] : 0.f;
          gaa[i][p] = (l < d.L && my_q >= 0) ? gq[2] : 0.f;
        }
      tile_front_backward<T>(front, d, b, m, my_q, h, a, off, sc, gxx, gaa);
    } else {
      tile_fallback_bwd<T, NU, 2, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, s_lvl, s_plan,
                                         gv_scale, s_q);
    }
    return;
  }
  float2 go[kWinQ][NU][E / 2];
#pragma unroll
  for (int q = 0; q < kWinQ; ++q)
#pragma unroll
    for (int uu = 0; uu < NU; ++uu) {
      float t[E];
      U::unpack_raw(gor[q][uu], t);
#pragma unroll
      for (int c = 0; c < E / 2; ++c) go[q][uu][c] = make_float2(t[2 * c], t[2 * c + 1]);
    }
  tile_stamp(4);
  if (pl.total > 0) tile_mbar_wait(&s_mbar, 0);
  tile_stamp(5);
  __shared__ unsigned char s_rowmask[FUSED ? kTileCapRows : 1];
  const bool has_mask = FUSED && front.mask != nullptr;
  if (has_mask)
    tile_apply_mask<RB>(front.mask + (size_t)b * d.S, d.S, pl.tlo, pl.tbase, pl.total, vtile, s_rowmask);

  const int MD = d.M * d.D;
  constexpr bool BULK = TileBulkRed<GV>::value;
  constexpr int GRB = NU * G * E * 4;   // bytes of one head's grad_value row
  char* s_stage = s_plan + kTileBlocks * kTileBlkRows * 16;
  int it = 0;                           // row pairs done (warp-uniform): the staging slot alternates
  GV* gvb = gvalue + ((size_t)b * d.S * d.M + m) * d.D + j * E;
  // warp-uniform row loops over pairs of rows: a block past its own rows re-reads its last row
  // against zero weights and issues no reds
#if MSDA_TILE_LEVEL_LOOP
#pragma unroll 1
#else
#pragma unroll
#endif
  for (int l = 0; l < kWinMaxL; ++l) {
    const int l_trow = tile_sel4(l, pl.trow), l_bpoff = tile_sel4(l, pl.bpoff);
    const char* vr = vtile + l_trow * RB + j * U::BYTES;
    const float4* wl = Wt + l_bpoff * kTileWtStride;
    float* gl = Gs + (l_bpoff + (j >> 2)) * (4 * kTileWtStride) + (j & 3);
    GV* gp = gvb + (size_t)tile_sel4(l, pl.blo) * MD;
    const int n = tile_sel4(l, pl.wrows), nb = tile_sel4(l, pl.brows);
    for (int i = 0; i < n; i += 2) {
      float x[2 * kWinQ];
      float* stg = nullptr;
      bool realr[2] = {false, false};
      GV* gp0 = gp;
      if (BULK) {
        // the slot was handed to the TMA unit two pairs ago: that group must have read it
        if (j == 0) tile_bulk_wait_read<1>();
        __syncwarp();
        stg = reinterpret_cast<float*>(s_stage + (size_t)(((it & 1) * kTileBlocks + g) * 2) * GRB) + j * E;
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const float4 w = wl[(i + r) * kTileWtStride];
        const float wq[kWinQ] = {w.x, w.y, w.z, w.w};
#ifdef MSDA_TILE_DBG_NORED
        const bool real = gv_scale == 12345.f;   // experiment: no reds
#else
        const bool real = i + r < nb && !(has_mask && s_rowmask[l_trow + i + r]);
#endif
        float2 s[kWinQ];
        // grad_value row: sum_q Wt[t][q] * grad_out[q]
        float2 dv[NU][E / 2];
#pragma unroll
        for (int u = 0; u < NU; ++u)
#pragma unroll
          for (int q = 0; q < kWinQ; ++q) {
            const float2 w2 = make_float2(wq[q], wq[q]);
#pragma unroll
            for (int c = 0; c < E / 2; ++c)
              dv[u][c] = q == 0 ? __fmul2_rn(w2, go[q][u][c]) : __ffma2_rn(w2, go[q][u][c], dv[u][c]);
          }
        if (BULK) {
#pragma unroll
          for (int u = 0; u < NU; ++u)
#pragma unroll
            for (int c = 0; c < E / 2; c += 2)
              *reinterpret_cast<float4*>(stg + r * (GRB / 4) + u * (G * E) + 2 * c) =
                  make_float4(dv[u][c].x, dv[u][c].y, dv[u][c + 1].x, dv[u][c + 1].y);
        } else if (MSDA_TILE_RED_GROUP) {
          // one branch around the row's reds (ptxas turns every predicated red into its own
          // BSSY / BRA / BSYNC region)
          if (real) {
#pragma unroll
            for (int u = 0; u < NU; ++u)
#pragma unroll
              for (int c = 0; c < E / 2; c += 2)
                GvRed<GV>::add4(gp + u * (G * E) + 2 * c, true, dv[u][c].x, dv[u][c].y, dv[u][c + 1].x,
                                dv[u][c + 1].y, gv_scale);
          }
        } else {
#pragma unroll
          for (int u = 0; u < NU; ++u)
#pragma unroll
            for (int c = 0; c < E / 2; c += 2)
              GvRed<GV>::add4(gp + u * (G * E) + 2 * c, real, dv[u][c].x, dv[u][c].y, dv[u][c + 1].x,
                              dv[u][c + 1].y, gv_scale);
        }
#ifdef MSDA_TILE_DBG_NODOTS
#pragma unroll
        for (int q = 0; q < kWinQ; ++q) s[q] = make_float2(wq[q], 0.f);   // experiment: no value rows, no dots
        if (false)
#endif
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          // this lane's share of <grad_out[q], value row>
          Raw raw[U::RW];
          tile_lds<U>(vr + u * (G * U::BYTES), raw);
          float v[E];
          U::unpack_raw(raw, v);
#pragma unroll
          for (int q = 0; q < kWinQ; ++q)
#pragma unroll
            for (int c = 0; c < E / 2; ++c) {
              const float2 v2 = make_float2(v[2 * c], v[2 * c + 1]);
              s[q] = (u == 0 && c == 0) ? __fmul2_rn(go[q][u][c], v2) : __ffma2_rn(go[q][u][c], v2, s[q]);
            }
        }
#pragma unroll
        for (int q = 0; q < kWinQ; ++q) x[r * kWinQ + q] = s[q].x + s[q].y;
        realr[r] = real;
        gp += MD;
        if (i + r + 1 < nb) vr += RB;
      }
#ifdef MSDA_TILE_DBG_NOTR
      const float tot = x[0] + x[1] + x[2] + x[3] + x[4] + x[5] + x[6] + x[7];   // experiment: no shuffles
#else
      const float tot = transpose_reduce<8>(x, j, 0xffffffffu);
#endif
      gl[i * (4 * kTileWtStride)] = tot;   // (row i + (j >> 2), query j & 3)
      if (BULK) {
        // the block's two rows, complete in shared memory -> one bulk reduction each, issued by
        // the block's first lane
        tile_fence_async_smem();
        __syncwarp();
        if (j == 0) {
          float* grow = reinterpret_cast<float*>(gp0);
#pragma unroll
          for (int r = 0; r < 2; ++r)
            if (realr[r]) tile_bulk_red_f32(grow + (size_t)r * MD, stg + r * (GRB / 4), GRB);
          tile_bulk_commit();
        }
        ++it;
      }
    }
  }
  if (BULK && j == 0) tile_bulk_wait_read<0>();   // shared memory must outlive the reads
  __syncwarp();
  tile_stamp(6);

  // per-point gradients: the points again, the geometry again, the dots
  const bool have_q = my_q >= 0;
  const int LP = d.L * kWinP;
  const size_t grp_me = ((size_t)b * d.Lq + (have_q ? my_q : 0)) * d.M + m;
  float fx[2][4], fy[2][4], fa[2][4], foff[2][4], fsc[2], fgx[2][4], fga[2][4];
  if (FUSED)
    tile_front_points<T>(front, d, grp_me, (size_t)b * d.Lq + (have_q ? my_q : 0), h, s_lvl, fx, fy, fa, foff, fsc);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int l = h + 2 * i;
    const bool active = have_q && l < d.L;
    const int ls = l < d.L ? l : 0;
    float x[4], y[4], a[4];
    if (FUSED) {
#pragma unroll
      for (int p = 0; p < 4; ++p) { x[p] = fx[i][p]; y[p] = fy[i][p]; a[p] = fa[i][p]; }
    } else {
      WinPts<T>::ld_stream(loc + (grp_me * LP + ls * kWinP) * 2, aw + grp_me * LP + ls * kWinP, x, y, a);
    }
    const int W = s_lvl[3 * ls + 1], start = s_lvl[3 * ls + 2];
    const int tb = i == 0 ? (h ? pl.bpoff[1] - pl.blo[1] : pl.bpoff[0] - pl.blo[0])
                          : (h ? pl.bpoff[3] - pl.blo[3] : pl.bpoff[2] - pl.blo[2]);
    float gx[4], gy[4], ga[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      int r0;
      float w0, w1;
      bool live, v0, v1;
      WinCoef c;
      win_point<true>(x[p], y[p], a[p], W, start, d.border != 0, active, r0, w0, w1, live, v0, v1, c);
      const int t0 = (r0 + tb) * (4 * kTileWtStride) + qi;
      const float g0 = v0 ? Gs[t0] : 0.f;
      const float g1 = v1 ? Gs[t0 + 4 * kTileWtStride] : 0.f;
      ga[p] = fmaf(c.k0, g0, c.k1 * g1);
      gx[p] = c.fx * (g1 - g0);
      gy[p] = fmaf(c.cy0, g0, c.cy1 * g1);
      if (FUSED) { fgx[i][p] = gx[p]; fga[i][p] = ga[p]; }
    }
    if (!FUSED && active)
      WinPts<T>::st(gloc + (grp_me * LP + l * kWinP) * 2, gaw + grp_me * LP + l * kWinP, gx, gy, ga);
  }
  if (FUSED) tile_front_backward<T>(front, d, b, m, my_q, h, fa, foff, fsc, fgx, fga);
  tile_stamp(7);
}

}  // namespace msda
