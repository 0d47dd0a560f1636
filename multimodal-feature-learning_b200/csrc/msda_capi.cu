// msda_capi.cu -- host side of libmsda_b200.so: argument validation, kernel
// selection and launches behind the C ABI declared in include/msda_b200.h.
// Replaces the reference host launchers ms_deform_attn_cuda_forward/backward
// (models/ops/src/cuda/ms_deform_attn_cuda.cu:20-153) and the D-dispatch table
// ms_deformable_col2im_cuda (ms_deform_im2col_cuda.cuh:957-1328).

#include <atomic>
#include <type_traits>

#include "../../include/msda_b200.h"
#include "msda_dam.cuh"
#include "msda_fast.cuh"
#include "msda_frontend.cuh"
#include "msda_generic.cuh"
#include "msda_plan.cuh"
#include "msda_prepare.cuh"
#include "msda_sparse.cuh"
#include "msda_tile.cuh"
#include "msda_win.cuh"

using namespace msda;

namespace {

thread_local int t_last_cuda_error = 0;

inline int cuda_fail(cudaError_t e) {
  t_last_cuda_error = (int)e;
  return MSDA_ERR_CUDA;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

inline int log2_exact(int64_t g) {  // -1 unless g in {1,2,4,8,16,32}
  for (int s = 0; s <= 5; ++s)
    if (g == (int64_t(1) << s)) return s;
  return -1;
}

// Lane mapping of the vectorised kernels: each of G lanes owns NU 16-byte units.
//   f32 : D in {64,128,256} -> NU=2 (8 channels/lane), D in {4,8,16,32} -> NU=1
//   bf16: D in {8,...,256}  -> NU=1 (8 channels/lane)
struct FastPlan { int nu, g; };
inline bool fast_plan(int64_t D, int dtype, FastPlan* p) {
  if (dtype == MSDA_F32) {
    if (D == 64 || D == 128 || D == 256) { p->nu = 2; p->g = (int)(D / 8); return true; }
    if (D == 4 || D == 8 || D == 16 || D == 32) { p->nu = 1; p->g = (int)(D / 4); return true; }
    return false;
  }
  if (dtype == MSDA_BF16 && D % 8 == 0 && log2_exact(D / 8) >= 0) {
    p->nu = 1; p->g = (int)(D / 8); return true;
  }
  return false;
}

inline bool fast_applies(int64_t D, int64_t L, int dtype, unsigned flags) {
  FastPlan p;
  if (flags & MSDA_FLAG_FORCE_GENERIC) return false;
  return L <= kMaxLevelsFast && fast_plan(D, dtype, &p);
}

int check_common(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P,
                 int dtype, int padding) {
  if (dtype != MSDA_F32 && dtype != MSDA_F64 && dtype != MSDA_BF16) return MSDA_ERR_BAD_DTYPE;
  if (padding != MSDA_PAD_ZEROS && padding != MSDA_PAD_BORDER) return MSDA_ERR_BAD_PADDING;
  if (N < 0 || S < 0 || Lq < 0 || M <= 0 || D <= 0 || L <= 0 || P <= 0) return MSDA_ERR_BAD_SHAPE;
  const int64_t lim = (int64_t(1) << 31) - 1;
  if (S > lim || N * Lq * M > lim || L * P > lim || M * D > lim) return MSDA_ERR_TOO_LARGE;
  return MSDA_OK;
}

inline Dims make_dims(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P,
                      int padding) {
  Dims d;
  d.N = (int)N; d.S = (int)S; d.M = (int)M; d.D = (int)D; d.L = (int)L; d.Lq = (int)Lq; d.P = (int)P;
  d.border = padding == MSDA_PAD_BORDER;
  return d;
}

// Queries per tile: a multiple k of the groups per CTA.  Larger tiles reuse
// tapped rows from L1; smaller ones fill the 148 SMs when N*M*Lq is small.
inline int pick_qtile(const Dims& d, int G, unsigned flags) {
  const int ng = kFastThreads / G;
  int k = (int)((flags >> 8) & 0xfffu);
  if (k == 0) {
    k = 4;
    while (k > 1 && (int64_t)d.N * d.M * ((d.Lq + ng * k - 1) / (ng * k)) < 148 * 8) k >>= 1;
  }
  return ng * k;
}

template <typename T, int NU, int G>
int launch_fwd_fast(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                    const void* aw, void* out, const Dims& d, unsigned flags, cudaStream_t st) {
  const int qtile = pick_qtile(d, G, flags);
  const int nqc = (d.Lq + qtile - 1) / qtile;
  const int64_t tiles = (int64_t)d.N * d.M * nqc;
  msda_fwd_fast_kernel<T, NU, G><<<(unsigned)tiles, kFastThreads, 0, st>>>(
      (const T*)value, shapes, lsi, (const T*)loc, (const T*)aw, (T*)out, d, qtile, nqc);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

template <typename T, int NU, int G, typename GV = float, typename U = Unit<T>>
int launch_bwd_fast(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                    const void* aw, const void* gout, GV* gvalue, void* gloc, void* gaw,
                    const Dims& d, unsigned flags, cudaStream_t st, const DetHeader* det = nullptr) {
  const int qtile = pick_qtile(d, G, flags);
  const int nqc = (d.Lq + qtile - 1) / qtile;
  const int64_t tiles = (int64_t)d.N * d.M * nqc;
  msda_bwd_fast_kernel<T, NU, G, GV, U><<<(unsigned)tiles, kFastThreads, 0, st>>>(
      (const T*)value, shapes, lsi, (const T*)loc, (const T*)aw, (const T*)gout, gvalue, (T*)gloc,
      (T*)gaw, d, qtile, nqc, det);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

// Row-window kernels (msda_win.cuh): temporal maps + query locality.  They are
// instantiated for P = 4 points, L <= 4 levels and 8 lanes per head row
// (f32: D = 64 or 32, bf16: D = 64); the kernels check H == 1 on the device and
// decide per block of 4 queries between the window and the per-tap body.
inline bool win_applies(const Dims& d, int dtype, unsigned flags) {
  if (flags & (MSDA_FLAG_NO_WINDOW | MSDA_FLAG_FORCE_GENERIC)) return false;
  // Query lists shorter than this are decoder-style (a few learned queries per
  // item, unrelated reference points): their 4-query blocks have no common window.
  if (d.Lq < kWinMinQueries && !(flags & MSDA_FLAG_FORCE_WINDOW)) return false;
  if (d.P != kWinP || d.L > kWinMaxL || d.S <= 0) return false;
  // the points of a query are staged with 16-byte cp.async pieces
  if ((d.L * d.P * (dtype == MSDA_BF16 ? 2 : 4)) % 16 != 0) return false;
  if ((int64_t)d.N * d.M * ((d.Lq + 7) / 8) >= (int64_t(1) << 31)) return false;
  // 32-bit BYTE offsets inside one batch item's slab (fp32 accumulation image included)
  if ((int64_t)d.S * d.M * d.D * 4 >= (int64_t(1) << 32)) return false;
  if (dtype == MSDA_F32) return d.D == 64 || d.D == 32;
  if (dtype == MSDA_BF16) return d.D == 64;
  return false;
}

// Persistent launch: as many CTAs as are resident at once; warp w of the grid
// takes units w, w + W, w + 2W, ...
template <typename K>
int win_grid(K kern, size_t smem, long long units, int* grid) {
  int dev = 0, sms = 0, per_sm = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e == cudaSuccess)
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kWinThreads, smem);
  if (e != cudaSuccess) return cuda_fail(e);
  if (per_sm < 1) per_sm = 1;
  const long long warps = kWinThreads / 32;
  long long g = (long long)sms * per_sm;
  if (g > (units + warps - 1) / warps) g = (units + warps - 1) / warps;
  *grid = (int)(g < 1 ? 1 : g);
  return MSDA_OK;
}

template <typename T, int NU, int G>
int launch_fwd_win(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                   const void* aw, void* out, const Dims& d, cudaStream_t st) {
  constexpr int qpu = (32 / G) * kWinQ;                 // queries per unit (one warp)
  const int nqc = (d.Lq + qpu - 1) / qpu;
  const long long units = (long long)d.N * d.M * nqc;
  constexpr size_t smem = (size_t)(kWinThreads / G) * WinSmem<T, 1, false, 0>::kGroupBytes;
  auto kern = msda_fwd_win_kernel<T, NU, G>;
  int grid = 1;
  const int rc = win_grid(kern, smem, units, &grid);
  if (rc != MSDA_OK) return rc;
  kern<<<grid, kWinThreads, smem, st>>>((const T*)value, shapes, lsi, (const T*)loc, (const T*)aw,
                                        (T*)out, d, nqc, (int)units);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

template <typename T, int NU, int G, typename GV = float, typename U = Unit<T>>
int launch_bwd_win(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                   const void* aw, const void* gout, GV* gvalue, void* gloc, void* gaw,
                   const Dims& d, cudaStream_t st, const DetHeader* det = nullptr) {
  constexpr int qpu = (32 / G) * kWinQ;                 // queries per unit (one warp)
  const int nqc = (d.Lq + qpu - 1) / qpu;
  const long long units = (long long)d.N * d.M * nqc;
  constexpr size_t smem = (size_t)(kWinThreads / G) * WinSmem<T, 1, false, 0>::kGroupBytes;
  auto kern = msda_bwd_win_kernel<T, NU, G, GV, U>;
  int grid = 1;
  const int rc = win_grid(kern, smem, units, &grid);
  if (rc != MSDA_OK) return rc;
  kern<<<grid, kWinThreads, smem, st>>>((const T*)value, shapes, lsi, (const T*)loc, (const T*)aw,
                                        (const T*)gout, gvalue, (T*)gloc, (T*)gaw, d, nqc, (int)units,
                                        det);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

// ---- CTA-tile kernels (msda_tile.cuh) -----------------------------------------
// The value tensor as a 2-D TMA tensor: [N*S rows][M*D channels], box = 8 rows x D channels of
// one head.  cuTensorMapEncodeTiled is resolved through the runtime (no link against libcuda);
// the descriptor is plain host arithmetic, rebuilt per call and passed as a kernel parameter,
// so the library keeps no state about the caller's buffers.
typedef CUresult (*TileEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);

TileEncodeFn tile_encoder() {
  static const TileEncodeFn fn = []() -> TileEncodeFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    const cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
      (void)cudaGetLastError();
      return nullptr;
    }
    return reinterpret_cast<TileEncodeFn>(p);
  }();
  return fn;
}

bool tile_value_map(CUtensorMap* map, const void* base, int dtype, const Dims& d) {
  const TileEncodeFn enc = tile_encoder();
  if (!enc) return false;
  const cuuint64_t esz = dtype == MSDA_BF16 ? 2 : 4;
  const cuuint64_t gdim[2] = {(cuuint64_t)d.M * d.D, (cuuint64_t)d.N * d.S};
  const cuuint64_t gstride[1] = {(cuuint64_t)d.M * d.D * esz};
  const cuuint32_t box[2] = {(cuuint32_t)d.D, (cuuint32_t)kTileBoxRows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, dtype == MSDA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                         2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Same shape class as the row-window kernels, plus what TMA wants.  The query order comes from a
// workspace (msda_forward_ex / msda_backward_ex); without one the row-window kernels run.
inline bool tile_applies(const Dims& d, int dtype, unsigned flags) {
  if (flags & MSDA_FLAG_NO_TILE) return false;
  if (!win_applies(d, dtype, flags)) return false;
  if ((int64_t)d.N * d.S >= (int64_t(1) << 31) || (int64_t)d.N * d.S < kTileBoxRows) return false;
  if ((int64_t)d.N * d.M * ((d.Lq + kTileQ - 1) / kTileQ) >= (int64_t(1) << 31)) return false;
  // Queries much sparser in time than the value rows (audio tokens over the video memory of the
  // multimodal encoder: 300 queries on 1200 rows): the 64 time-adjacent queries of a tile would
  // span ~ 64 * S / Lq (+ ~6 per level) rows, beyond the tile, and every tile would run the per-tap
  // body.  The row-window kernels, whose unit is the 4-query block, take such calls (measured at
  // MM_AV: step 0.225 of the roofline against 0.166).  MSDA_FLAG_FORCE_WINDOW keeps the tiles.
  if (!(flags & MSDA_FLAG_FORCE_WINDOW) &&
      (int64_t)kTileQ * d.S > (int64_t)(kTileCapRows - 6 * d.L) * d.Lq)
    return false;
  return true;
}

inline size_t tile_plan_bytes(int64_t N, int64_t Lq) {
  return ((size_t)N * Lq * sizeof(int) + 255) & ~size_t(255);
}

// opt in to > 48 KB of dynamic shared memory, once per kernel and device
template <typename K>
int tile_smem_attr(K kern, size_t smem, std::atomic<unsigned long long>& done) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e);
  const unsigned long long bit = 1ull << (dev & 63);
  if (dev < 64 && (done.load(std::memory_order_acquire) & bit)) return MSDA_OK;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_fail(e);
  if (dev < 64) done.fetch_or(bit, std::memory_order_release);
  return MSDA_OK;
}

template <typename T>
int launch_sort(const SortSrc<T>& src, int* perm, const Dims& d, cudaStream_t st) {
  int hk_shift = 0;
  // heads averaged into the sort key: 2 (a key costs one fully divergent 32-byte sector per head)
  while (hk_shift < 1 && (2 << hk_shift) <= d.M) ++hk_shift;
  if (d.Lq <= kSortStash)   // cluster sort: deterministic order, eight SMs per batch item on the keys
    msda_sort_cluster_kernel<T><<<(unsigned)d.N * kSortCluster, kSortThreads, 0, st>>>(src, perm, d, hk_shift);
  else
    msda_sort_kernel<T><<<(unsigned)d.N, kSortThreads, 0, st>>>(src, perm, d, hk_shift);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}
int launch_sort_dtype(int dtype, const void* loc, int* perm, const Dims& d, cudaStream_t st) {
  if (dtype == MSDA_BF16) {
    const SortSrc<__nv_bfloat16> src = {(const __nv_bfloat16*)loc, nullptr, nullptr, nullptr, 0};
    return launch_sort<__nv_bfloat16>(src, perm, d, st);
  }
  const SortSrc<float> src = {(const float*)loc, nullptr, nullptr, nullptr, 0};
  return launch_sort<float>(src, perm, d, st);
}

// Launch a tile kernel as a programmatic dependent of the kernel before it in the stream (the
// sort, when this call computed the plan): its CTAs are scheduled while that kernel drains and
// wait in griddepcontrol.wait before they read the plan.
template <typename... KArgs, typename... Args>
int launch_pdl(void (*kern)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st,
               bool dependent, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = dependent ? 1 : 0;
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

template <typename T, int NU, bool FUSED = false>
int launch_fwd_tile(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                    const void* aw, const int* perm, void* out, const Dims& d, int dtype,
                    cudaStream_t st, bool after_sort, TileFront<T> front = TileFront<T>(),
                    char* save = nullptr) {
  static std::atomic<unsigned long long> attr_done{0};
  CUtensorMap map;
  if (!tile_value_map(&map, value, dtype, d)) return MSDA_ERR_NO_TMA;
  auto kern = msda_fwd_tile_kernel<T, NU, FUSED>;
  constexpr size_t smem = TileSmem<T, NU>::kBytes;
  const int rc = tile_smem_attr(kern, smem, attr_done);
  if (rc != MSDA_OK) return rc;
  const int nchunks = (d.Lq + kTileQ - 1) / kTileQ;
  const unsigned grid = (unsigned)((int64_t)d.N * d.M * nchunks);
  return launch_pdl(kern, grid, kTileThreads, smem, st, after_sort, map, (const T*)value, shapes, lsi,
                    (const T*)loc, (const T*)aw, perm, (T*)out, d, nchunks, front, save);
}

template <typename T, int NU, typename GV = float, typename U = Unit<T>, bool FUSED = false>
int launch_bwd_tile(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                    const void* aw, const void* gout, const int* perm, GV* gvalue, void* gloc,
                    void* gaw, const Dims& d, int dtype, cudaStream_t st,
                    const DetHeader* det = nullptr, bool after_sort = false,
                    TileFront<T> front = TileFront<T>(), const char* saved = nullptr) {
  static std::atomic<unsigned long long> attr_done{0};
  CUtensorMap map;
  if (!tile_value_map(&map, value, dtype, d)) return MSDA_ERR_NO_TMA;
  auto kern = msda_bwd_tile_kernel<T, NU, GV, U, FUSED>;
  constexpr size_t smem = tile_bwd_smem<T, NU, GV, U>();
  const int rc = tile_smem_attr(kern, smem, attr_done);
  if (rc != MSDA_OK) return rc;
  const int nchunks = (d.Lq + kTileQ - 1) / kTileQ;
  const unsigned grid = (unsigned)((int64_t)d.N * d.M * nchunks);
  return launch_pdl(kern, grid, kTileThreads, smem, st, after_sort, map, (const T*)value, shapes, lsi,
                    (const T*)loc, (const T*)aw, (const T*)gout, perm, gvalue, (T*)gloc, (T*)gaw, d,
                    nchunks, det, front, saved);
}

// ---- plan / apply split (msda_plan.cuh) ---------------------------------------------
// workspace of the split path: [query order: int32 N*Lq, 256-byte rounded][plan: tiles * kPlanStride]
inline size_t plan_tiles(int64_t N, int64_t M, int64_t Lq) {
  return (size_t)N * M * ((Lq + kTileQ - 1) / kTileQ);
}
inline size_t plan_ws_bytes(int64_t N, int64_t M, int64_t Lq) {
  return tile_plan_bytes(N, Lq) + plan_tiles(N, M, Lq) * kPlanStride;
}
// workspace with SAVED plans (MSDA_FLAG_SAVE_PLAN): [query order][tiles * kSaveStride]
inline size_t save_ws_bytes(int64_t N, int64_t M, int64_t Lq) {
  return tile_plan_bytes(N, Lq) + plan_tiles(N, M, Lq) * (size_t)kSaveStride;
}
inline bool save_applies(unsigned flags) {
  return MSDA_TILE_SAVED_PLANS != 0 && MSDA_TILE_WT_INTERLEAVE != 0 && (flags & MSDA_FLAG_SAVE_PLAN) &&
         !(flags & MSDA_FLAG_PLAN);
}
inline char* plan_of(void* workspace, int64_t N, int64_t Lq) {
  return (char*)workspace + tile_plan_bytes(N, Lq);
}

template <typename T>
int launch_plan(const int64_t* shapes, const int64_t* lsi, const void* loc, const void* aw,
                const int* perm, char* plan, const Dims& d, cudaStream_t st) {
  const int nchunks = (d.Lq + kTileQ - 1) / kTileQ;
  const unsigned grid = (unsigned)((int64_t)d.N * d.M * nchunks);
  return launch_pdl(msda_plan_kernel<T>, grid, kTileThreads, 0, st, true, shapes, lsi, (const T*)loc,
                    (const T*)aw, perm, plan, d, nchunks);
}

// persistent form of the forward tile kernel (fp32): one CTA per resident slot, walking tiles
template <int NU>
int launch_fwd_tile_persist(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                            const void* aw, const int* perm, void* out, const Dims& d, int dtype,
                            cudaStream_t st, bool after_sort) {
  static std::atomic<unsigned long long> attr_done{0};
  CUtensorMap map;
  if (!tile_value_map(&map, value, dtype, d)) return MSDA_ERR_NO_TMA;
  auto kern = msda_fwd_tile_persist_kernel<NU>;
  constexpr size_t smem = TileSmem<float, NU>::kBytes;
  const int rc = tile_smem_attr(kern, smem, attr_done);
  if (rc != MSDA_OK) return rc;
  const int nchunks = (d.Lq + kTileQ - 1) / kTileQ;
  const long long tiles = (long long)d.N * d.M * nchunks;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long slots = (long long)sms * MSDA_TILE_FWD_CTAS;
  return launch_pdl(kern, (unsigned)(tiles < slots ? tiles : slots), kTileThreads, smem, st, after_sort, map,
                    (const float*)value, shapes, lsi, (const float*)loc, (const float*)aw, perm, (float*)out, d,
                    nchunks, (int)tiles);
}

// persistent apply kernels: two CTAs per SM
inline int apply_grid(long long tiles) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long g = 2ll * sms;
  return (int)(tiles < g ? tiles : g);
}

template <typename T, int NU>
int launch_fwd_apply(const void* value, const int64_t* shapes, const int64_t* lsi, const void* loc,
                     const void* aw, const char* plan, void* out, const Dims& d, int dtype,
                     cudaStream_t st, bool after_plan) {
  static std::atomic<unsigned long long> attr_done{0};
  CUtensorMap map;
  if (!tile_value_map(&map, value, dtype, d)) return MSDA_ERR_NO_TMA;
  auto kern = msda_fwd_apply_kernel<T, NU>;
  constexpr size_t smem = ApplySmem<NU * 8 * 16>::kBytes;
  const int rc = tile_smem_attr(kern, smem, attr_done);
  if (rc != MSDA_OK) return rc;
  const int nchunks = (d.Lq + kTileQ - 1) / kTileQ;
  const long long tiles = (long long)d.N * d.M * nchunks;
  return launch_pdl(kern, (unsigned)apply_grid(tiles), kApplyThreads, smem, st, after_plan, map,
                    (const T*)value, shapes, lsi, (const T*)loc, (const T*)aw, plan, (T*)out, d, nchunks,
                    (int)tiles);
}

// (dtype, NU, G) -> instantiation.  FN is launch_fwd_fast / launch_bwd_fast.
#define MSDA_CASE(T, NU_, G_, FN, ...) \
  if (plan.nu == NU_ && plan.g == G_) return FN<T, NU_, G_>(__VA_ARGS__)
#define MSDA_DISPATCH_FAST(DTYPE, FN, ...)                       \
  do {                                                           \
    if ((DTYPE) == MSDA_F32) {                                   \
      MSDA_CASE(float, 2, 8, FN, __VA_ARGS__);                   \
      MSDA_CASE(float, 2, 16, FN, __VA_ARGS__);                  \
      MSDA_CASE(float, 2, 32, FN, __VA_ARGS__);                  \
      MSDA_CASE(float, 1, 1, FN, __VA_ARGS__);                   \
      MSDA_CASE(float, 1, 2, FN, __VA_ARGS__);                   \
      MSDA_CASE(float, 1, 4, FN, __VA_ARGS__);                   \
      MSDA_CASE(float, 1, 8, FN, __VA_ARGS__);                   \
    } else {                                                     \
      MSDA_CASE(__nv_bfloat16, 1, 1, FN, __VA_ARGS__);           \
      MSDA_CASE(__nv_bfloat16, 1, 2, FN, __VA_ARGS__);           \
      MSDA_CASE(__nv_bfloat16, 1, 4, FN, __VA_ARGS__);           \
      MSDA_CASE(__nv_bfloat16, 1, 8, FN, __VA_ARGS__);           \
      MSDA_CASE(__nv_bfloat16, 1, 16, FN, __VA_ARGS__);          \
      MSDA_CASE(__nv_bfloat16, 1, 32, FN, __VA_ARGS__);          \
    }                                                            \
    return MSDA_ERR_BAD_SHAPE;                                   \
  } while (0)

template <typename T>
int launch_fwd_generic(const void* value, const int64_t* shapes, const int64_t* lsi,
                       const void* loc, const void* aw, void* out, const Dims& d, cudaStream_t st) {
  const size_t total = (size_t)d.N * d.Lq * d.M * d.D;
  size_t blocks = (total + 255) / 256;
  if (blocks > 148u * 64u) blocks = 148u * 64u;
  msda_fwd_generic_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(
      (const T*)value, shapes, lsi, (const T*)loc, (const T*)aw, (T*)out, d, total);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

template <typename T, typename GV>
int launch_bwd_generic(const void* value, const int64_t* shapes, const int64_t* lsi,
                       const void* loc, const void* aw, const void* gout, GV* gvalue, void* gloc,
                       void* gaw, const Dims& d, cudaStream_t st, const DetHeader* det = nullptr) {
  const int64_t groups = (int64_t)d.N * d.Lq * d.M;
  msda_bwd_generic_kernel<T, GV><<<(unsigned)groups, 128, 0, st>>>(
      (const T*)value, shapes, lsi, (const T*)loc, (const T*)aw, (const T*)gout, gvalue, (T*)gloc,
      (T*)gaw, d, det);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}


// ---- deterministic backward (msda_det.cuh) ------------------------------------
// family: 3 = CTA-tile kernels (perm = the time order of the queries), 2 = row-window kernels,
// 1 = per-tap vectorised kernels (all only in their 8-lane instantiations: f32 D = 64 / 32,
// bf16 D = 64), 0 = any-D kernels
template <typename T, int NU>
int backward_deterministic(const void* value, const int64_t* shapes, const int64_t* lsi,
                           const void* loc, const void* aw, const void* gout, void* gvalue,
                           void* gloc, void* gaw, const Dims& d, void* workspace, int family,
                           unsigned flags, cudaStream_t st, const int* perm, int dtype) {
  DetHeader* hdr = (DetHeader*)workspace;
  long long* acc = (long long*)((char*)workspace + sizeof(DetHeader));
  const size_t nvalue = (size_t)d.N * d.S * d.M * d.D;
  const size_t ngo = (size_t)d.N * d.Lq * d.M * d.D, naw = (size_t)d.N * d.Lq * d.M * d.L * d.P;
  auto blocks = [](size_t n) { size_t b = (n + 255) / 256; return (unsigned)(b > 148u * 8u ? 148u * 8u : b < 1 ? 1 : b); };
  msda_det_amax_kernel<T><<<blocks(ngo), 256, 0, st>>>((const T*)gout, ngo, &hdr->amax_go);
  msda_det_amax_kernel<T><<<blocks(naw), 256, 0, st>>>((const T*)aw, naw, &hdr->amax_aw);
  int log2_addends = 0;
  while ((int64_t(1) << log2_addends) < (int64_t)d.Lq * d.L * d.P * 4) ++log2_addends;
  msda_det_scale_kernel<<<1, 1, 0, st>>>(hdr, log2_addends);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e);
  int r = MSDA_OK;
  bool done = false;
  if constexpr (!std::is_same<T, double>::value) {   // the vectorised kernels have no fp64 form
    if (family == 3) {
      r = launch_bwd_tile<T, NU, long long>(value, shapes, lsi, loc, aw, gout, perm, acc, gloc, gaw, d, dtype, st, hdr);
      done = true;
    } else if (family == 2) {
      r = launch_bwd_win<T, NU, 8, long long>(value, shapes, lsi, loc, aw, gout, acc, gloc, gaw, d, st, hdr);
      done = true;
    } else if (family == 1) {
      r = launch_bwd_fast<T, NU, 8, long long>(value, shapes, lsi, loc, aw, gout, acc, gloc, gaw, d, flags, st, hdr);
      done = true;
    }
  }
  if (!done)
    r = launch_bwd_generic<T, long long>(value, shapes, lsi, loc, aw, gout, acc, gloc, gaw, d, st, hdr);
  if (r != MSDA_OK) return r;
  msda_det_convert_kernel<T><<<blocks(nvalue), 256, 0, st>>>(acc, (T*)gvalue, nvalue, hdr);
  e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

// ---- module front-end (msda_frontend.cuh) -----------------------------------
int check_front(int64_t N, int64_t Lq, int64_t M, int64_t L, int64_t P, int ref_dim, int dtype) {
  if (N < 0 || Lq < 0 || M <= 0 || L <= 0 || P <= 0) return MSDA_ERR_BAD_SHAPE;
  if (ref_dim != 1 && ref_dim != 2) return MSDA_ERR_BAD_SHAPE;
  if (dtype != MSDA_F32 && dtype != MSDA_F64 && dtype != MSDA_BF16) return MSDA_ERR_BAD_DTYPE;
  if (L * P >= (int64_t(1) << 20) || M >= (int64_t(1) << 20)) return MSDA_ERR_TOO_LARGE;
  if (N * Lq >= (int64_t(1) << 40) / M) return MSDA_ERR_TOO_LARGE;
  return MSDA_OK;
}

template <typename T>
int launch_front_fwd(const void* off, const void* logits, const void* ref, const int64_t* tl,
                     void* loc, void* aw, const FrontDims& d, cudaStream_t st) {
  const unsigned grid = (unsigned)((d.groups + kFrontThreads - 1) / kFrontThreads);
  // 4 points per 8/16/32-byte piece when P allows it and the arrays are aligned for it
  const bool vec = d.P % 4 == 0 && aligned16(off) && aligned16(logits) && aligned16(loc) && aligned16(aw);
  const int tpg = (int)(d.L * d.P / 4);          // 4-element pieces per group
  if (vec && log2_exact(tpg) >= 0) {
    const unsigned lgrid = (unsigned)((d.groups * tpg + kFrontThreads - 1) / kFrontThreads);
    msda_frontend_fwd_lanes_kernel<T><<<lgrid, kFrontThreads, 0, st>>>(
        (const T*)off, (const T*)logits, (const T*)ref, tl, (T*)loc, (T*)aw, d, tpg);
  } else if (vec)
    msda_frontend_fwd_kernel<T, 4><<<grid, kFrontThreads, 0, st>>>(
        (const T*)off, (const T*)logits, (const T*)ref, tl, (T*)loc, (T*)aw, d);
  else
    msda_frontend_fwd_kernel<T, 1><<<grid, kFrontThreads, 0, st>>>(
        (const T*)off, (const T*)logits, (const T*)ref, tl, (T*)loc, (T*)aw, d);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

template <typename T>
int launch_front_bwd(const void* off, const void* aw, const void* ref, const int64_t* tl,
                     const void* gloc, const void* gaw, void* goff, void* glogits, void* gref,
                     int64_t gref_elems, const FrontDims& d, cudaStream_t st) {
  const unsigned grid = (unsigned)((d.groups + kFrontThreads - 1) / kFrontThreads);
  const bool vec = d.P % 4 == 0 && aligned16(off) && aligned16(aw) && aligned16(gloc) &&
                   aligned16(gaw) && aligned16(goff) && aligned16(glogits);
  const int tpg = (int)(d.L * d.P / 4);
  const bool lanes = vec && log2_exact(tpg) >= 0;
  // shuffles reduce grad_ref over the heads of a query when those share a warp
  const bool shfl = log2_exact(d.M) >= 0 && (!lanes || d.M * tpg <= 32);
  if (gref && !shfl) {   // atomics accumulate into zeros
    const cudaError_t e = cudaMemsetAsync(gref, 0, (size_t)gref_elems * sizeof(T), st);
    if (e != cudaSuccess) return cuda_fail(e);
  }
#define MSDA_FRONT_BWD(V_, S_)                                                              \
  msda_frontend_bwd_kernel<T, V_, S_><<<grid, kFrontThreads, 0, st>>>(                       \
      (const T*)off, (const T*)aw, (const T*)ref, tl, (const T*)gloc, (const T*)gaw, (T*)goff, \
      (T*)glogits, (T*)gref, d)
  if (lanes) {
    const unsigned lgrid = (unsigned)((d.groups * tpg + kFrontThreads - 1) / kFrontThreads);
    if (shfl)
      msda_frontend_bwd_lanes_kernel<T, true><<<lgrid, kFrontThreads, 0, st>>>(
          (const T*)off, (const T*)aw, (const T*)ref, tl, (const T*)gloc, (const T*)gaw, (T*)goff,
          (T*)glogits, (T*)gref, d, tpg);
    else
      msda_frontend_bwd_lanes_kernel<T, false><<<lgrid, kFrontThreads, 0, st>>>(
          (const T*)off, (const T*)aw, (const T*)ref, tl, (const T*)gloc, (const T*)gaw, (T*)goff,
          (T*)glogits, (T*)gref, d, tpg);
  } else if (vec && shfl) MSDA_FRONT_BWD(4, true);
  else if (vec) MSDA_FRONT_BWD(4, false);
  else if (shfl) MSDA_FRONT_BWD(1, true);
  else MSDA_FRONT_BWD(1, false);
#undef MSDA_FRONT_BWD
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

}  // namespace

extern "C" {

int msda_abi_version(void) { return MSDA_B200_ABI_VERSION; }

const char* msda_error_string(int status) {
  switch (status) {
    case MSDA_OK: return "ok";
    case MSDA_ERR_NULL_POINTER: return "null pointer argument";
    case MSDA_ERR_BAD_SHAPE: return "invalid shape (M, D, L, P must be > 0; N, S, Lq >= 0)";
    case MSDA_ERR_BAD_DTYPE: return "unsupported dtype (expected f32, f64 or bf16)";
    case MSDA_ERR_BAD_PADDING: return "unsupported padding mode (expected zeros or border)";
    case MSDA_ERR_TOO_LARGE: return "problem exceeds the 2^31-1 row/group index range";
    case MSDA_ERR_WORKSPACE: return "workspace missing or smaller than msda_backward_workspace_bytes()";
    case MSDA_ERR_CUDA: return "CUDA runtime error (see msda_last_cuda_error_string)";
    case MSDA_ERR_MISALIGNED: return "data pointer not aligned to its element size";
    case MSDA_ERR_UNSUPPORTED: return "shape / dtype / flags outside the fused module path (use msda_frontend_* + msda_forward/backward)";
    case MSDA_ERR_SCRATCH_LIMIT: return "msda_forward_static: the call needs more than the built-in MSDA_STATIC_SCRATCH_BYTES of scratch (N * Lq > 262144); use msda_forward_ex with a workspace";
    case MSDA_ERR_NO_TMA: return "TMA tensor map could not be encoded (driver entry point missing or rejected)";
    default: return "unknown msda status";
  }
}

const char* msda_kernel_family(int64_t D, int dtype, unsigned flags) {
  return fast_applies(D, 1, dtype, flags) ? "fast" : "generic";
}

const char* msda_kernel_choice(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq,
                               int64_t P, int dtype, unsigned flags) {
  if (check_common(N, S, M, D, L, Lq, P, dtype, MSDA_PAD_ZEROS) != MSDA_OK) return "invalid";
  const bool fast = S * M * D < (int64_t(1) << 31) && fast_applies(D, L, dtype, flags);
  if (!fast) return "generic";
  const Dims d = make_dims(N, S, M, D, L, Lq, P, MSDA_PAD_ZEROS);
  if (tile_applies(d, dtype, flags)) return "tile";
  return win_applies(d, dtype, flags) ? "window" : "fast";
}

#ifdef MSDA_TILE_TIMING
int msda_debug_tile_times(unsigned long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_tile_times, (size_t)n * sizeof(unsigned long long));
}
int msda_debug_sort_times(unsigned long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_sort_times, (size_t)n * sizeof(unsigned long long));
}
#endif

int msda_last_cuda_error(void) { return t_last_cuda_error; }
const char* msda_last_cuda_error_string(void) {
  return cudaGetErrorString((cudaError_t)t_last_cuda_error);
}

size_t msda_forward_workspace_bytes(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                                    int64_t Lq, int64_t P, int dtype, unsigned flags) {
  if (check_common(N, S, M, D, L, Lq, P, dtype, MSDA_PAD_ZEROS) != MSDA_OK || N == 0 || Lq == 0) return 0;
  const bool fast = S * M * D < (int64_t(1) << 31) && fast_applies(D, L, dtype, flags);
  if (!fast || !tile_applies(make_dims(N, S, M, D, L, Lq, P, MSDA_PAD_ZEROS), dtype, flags)) return 0;
  if (flags & MSDA_FLAG_PLAN) return plan_ws_bytes(N, M, Lq);
  return save_applies(flags) ? save_ws_bytes(N, M, Lq) : tile_plan_bytes(N, Lq);
}

int msda_forward_ex(const void* value, const int64_t* spatial_shapes, const int64_t* level_start_index,
                    const void* sampling_loc, const void* attn_weight, void* output, int64_t N,
                    int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P, int dtype,
                    int padding, unsigned flags, void* workspace, size_t workspace_bytes,
                    void* stream) {
  const int rc = check_common(N, S, M, D, L, Lq, P, dtype, padding);
  if (rc != MSDA_OK) return rc;
  if (N == 0 || Lq == 0) return MSDA_OK;  // empty output
  if (!spatial_shapes || !level_start_index || !sampling_loc || !attn_weight || !output)
    return MSDA_ERR_NULL_POINTER;
  if (S > 0 && !value) return MSDA_ERR_NULL_POINTER;
  Dims d = make_dims(N, S, M, D, L, Lq, P, padding);
  cudaStream_t st = (cudaStream_t)stream;
  const bool fits32 = S * M * D < (int64_t(1) << 31);
  const bool fast = fits32 && fast_applies(D, L, dtype, flags) && aligned16(value) && aligned16(output) &&
                    aligned16(sampling_loc) && aligned16(attn_weight);
  // CTA tiles in time order: needs the plan workspace (the query permutation) and TMA
  if (fast && tile_applies(d, dtype, flags) && workspace && aligned16(workspace) &&
      workspace_bytes >= tile_plan_bytes(N, Lq) && tile_encoder()) {
    int* perm = (int*)workspace;
    int r = launch_sort_dtype(dtype, sampling_loc, perm, d, st);
    if (r != MSDA_OK) return r;
    if ((flags & MSDA_FLAG_PLAN) && workspace_bytes >= plan_ws_bytes(N, M, Lq) &&
        (tile_plan_bytes(N, Lq) % 128) == 0 && (reinterpret_cast<uintptr_t>(workspace) & 127u) == 0) {
      // plan kernel + apply kernel
      char* plan = plan_of(workspace, N, Lq);
      r = dtype == MSDA_BF16
              ? launch_plan<__nv_bfloat16>(spatial_shapes, level_start_index, sampling_loc, attn_weight, perm, plan, d, st)
              : launch_plan<float>(spatial_shapes, level_start_index, sampling_loc, attn_weight, perm, plan, d, st);
      if (r != MSDA_OK) return r;
      if (dtype == MSDA_BF16)
        r = launch_fwd_apply<__nv_bfloat16, 1>(value, spatial_shapes, level_start_index, sampling_loc,
                                               attn_weight, plan, output, d, dtype, st, true);
      else if (D == 64)
        r = launch_fwd_apply<float, 2>(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                                       plan, output, d, dtype, st, true);
      else
        r = launch_fwd_apply<float, 1>(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                                       plan, output, d, dtype, st, true);
      if (r != MSDA_ERR_NO_TMA) return r;
    }
    // MSDA_FLAG_SAVE_PLAN: the tiles' plans go behind the query order, for msda_backward_ex
    char* save = save_applies(flags) && workspace_bytes >= save_ws_bytes(N, M, Lq) &&
                         (reinterpret_cast<uintptr_t>(workspace) & 127u) == 0
                     ? (char*)workspace + tile_plan_bytes(N, Lq)
                     : nullptr;
    if ((flags & MSDA_FLAG_PERSIST) && dtype == MSDA_F32 && !save)
      r = D == 64 ? launch_fwd_tile_persist<2>(value, spatial_shapes, level_start_index, sampling_loc,
                                               attn_weight, perm, output, d, dtype, st, true)
                  : launch_fwd_tile_persist<1>(value, spatial_shapes, level_start_index, sampling_loc,
                                               attn_weight, perm, output, d, dtype, st, true);
    else if (dtype == MSDA_BF16)
      r = launch_fwd_tile<__nv_bfloat16, 1>(value, spatial_shapes, level_start_index, sampling_loc,
                                            attn_weight, perm, output, d, dtype, st, true,
                                            TileFront<__nv_bfloat16>(), save);
    else if (D == 64)
      r = launch_fwd_tile<float, 2>(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                                    perm, output, d, dtype, st, true, TileFront<float>(), save);
    else
      r = launch_fwd_tile<float, 1>(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                                    perm, output, d, dtype, st, true, TileFront<float>(), save);
    if (r != MSDA_ERR_NO_TMA) return r;   // descriptor rejected: the row-window kernels take over
  }
  if (fast && win_applies(d, dtype, flags)) {
    if (dtype == MSDA_BF16)
      return launch_fwd_win<__nv_bfloat16, 1, 8>(value, spatial_shapes, level_start_index,
                                                 sampling_loc, attn_weight, output, d, st);
    if (D == 64)
      return launch_fwd_win<float, 2, 8>(value, spatial_shapes, level_start_index, sampling_loc,
                                         attn_weight, output, d, st);
    return launch_fwd_win<float, 1, 8>(value, spatial_shapes, level_start_index, sampling_loc,
                                       attn_weight, output, d, st);
  }
  if (fast) {
    FastPlan plan;
    fast_plan(D, dtype, &plan);
    MSDA_DISPATCH_FAST(dtype, launch_fwd_fast, value, spatial_shapes, level_start_index,
                       sampling_loc, attn_weight, output, d, flags, st);
  }
  switch (dtype) {
    case MSDA_F32:
      return launch_fwd_generic<float>(value, spatial_shapes, level_start_index, sampling_loc,
                                       attn_weight, output, d, st);
    case MSDA_F64:
      return launch_fwd_generic<double>(value, spatial_shapes, level_start_index, sampling_loc,
                                        attn_weight, output, d, st);
    default:
      return launch_fwd_generic<__nv_bfloat16>(value, spatial_shapes, level_start_index,
                                               sampling_loc, attn_weight, output, d, st);
  }
}

int msda_forward(const void* value, const int64_t* spatial_shapes, const int64_t* level_start_index,
                 const void* sampling_loc, const void* attn_weight, void* output, int64_t N,
                 int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P, int dtype,
                 int padding, unsigned flags, void* stream) {
  return msda_forward_ex(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, output,
                         N, S, M, D, L, Lq, P, dtype, padding, flags, nullptr, 0, stream);
}

}  // extern "C"

// ---- msda_forward_static: the plan workspace as a built-in device array -------------
// No counterpart in the reference (added for callers
// that cannot allocate).  One array per GPU that loads the library, shared by every call on
// that GPU; it keeps the last call's plan.  Nothing in it has to start at zero: the sort
// kernel writes all N * Lq entries of the plan before the tile kernel reads them.
static __device__ __align__(256) unsigned char g_msda_static_scratch[MSDA_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::atomic<void*> g_static_scratch_ptr[kStaticMaxDevices];

// Device address of the array on the CURRENT device.  In host code the array's name is a
// host-side stand-in, not a device address: resolved once per device with
// cudaGetSymbolAddress (while that device is current) and cached.
int static_scratch(void** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e);
  if (dev < 0 || dev >= kStaticMaxDevices) return MSDA_ERR_SCRATCH_LIMIT;
  void* p = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    e = cudaGetSymbolAddress(&p, g_msda_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e);
    if (!p) return MSDA_ERR_CUDA;
    g_static_scratch_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return MSDA_OK;
}
}  // namespace

extern "C" int msda_forward_static(const void* value, const int64_t* spatial_shapes,
                                   const int64_t* level_start_index, const void* sampling_loc,
                                   const void* attn_weight, void* output, int64_t N, int64_t S, int64_t M,
                                   int64_t D, int64_t L, int64_t Lq, int64_t P, int dtype, int padding,
                                   unsigned flags, void* stream) {
  // the plan / apply split and saved plans keep ~14 KB per tile: not from this array
  flags &= ~(MSDA_FLAG_PLAN | MSDA_FLAG_SAVE_PLAN);
  const int rc = check_common(N, S, M, D, L, Lq, P, dtype, padding);
  if (rc != MSDA_OK) return rc;
  const size_t need = msda_forward_workspace_bytes(N, S, M, D, L, Lq, P, dtype, flags);
  if (need > (size_t)MSDA_STATIC_SCRATCH_BYTES) return MSDA_ERR_SCRATCH_LIMIT;
  void* scratch = nullptr;
  if (need > 0) {
    const int r = static_scratch(&scratch);
    if (r != MSDA_OK) return r;
  }
  return msda_forward_ex(value, spatial_shapes, level_start_index, sampling_loc, attn_weight, output, N, S,
                         M, D, L, Lq, P, dtype, padding, flags, scratch,
                         need > 0 ? (size_t)MSDA_STATIC_SCRATCH_BYTES : 0, stream);
}

namespace {
// the part of the backward workspace that is REQUIRED: the fp32 accumulation image of a bf16
// grad_value, or the fixed-point image + header of the deterministic variant
size_t ws_main_bytes(int64_t N, int64_t S, int64_t M, int64_t D, int dtype, unsigned flags) {
  if (N <= 0 || S <= 0 || M <= 0 || D <= 0) return 0;
  if (flags & MSDA_FLAG_DETERMINISTIC) return sizeof(DetHeader) + (size_t)N * S * M * D * sizeof(long long);
  if (dtype != MSDA_BF16) return 0;
  return (size_t)N * S * M * D * sizeof(float);
}
inline size_t round256(size_t x) { return (x + 255) & ~size_t(255); }
}  // namespace

extern "C" {

size_t msda_backward_workspace_bytes_flags(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                                           int64_t Lq, int64_t P, int dtype, unsigned flags) {
  const size_t main_bytes = ws_main_bytes(N, S, M, D, dtype, flags);
  // + the plan of the CTA-tile kernels (optional: without it the row-window kernels run)
  const size_t plan = msda_forward_workspace_bytes(N, S, M, D, L, Lq, P, dtype,
                                                   flags & ~(MSDA_FLAG_PLAN | MSDA_FLAG_SAVE_PLAN));
  return plan ? round256(main_bytes) + plan : main_bytes;
}

size_t msda_backward_workspace_bytes(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L,
                                     int64_t Lq, int64_t P, int dtype) {
  return msda_backward_workspace_bytes_flags(N, S, M, D, L, Lq, P, dtype, 0);
}

int msda_backward_ex(const void* value, const int64_t* spatial_shapes, const int64_t* level_start_index,
                     const void* sampling_loc, const void* attn_weight, const void* grad_output,
                     void* grad_value, void* grad_sampling_loc, void* grad_attn_weight, int64_t N,
                     int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P, int dtype,
                     int padding, unsigned flags, void* workspace, size_t workspace_bytes,
                     const void* plan, void* stream) {
  const int rc = check_common(N, S, M, D, L, Lq, P, dtype, padding);
  if (rc != MSDA_OK) return rc;
  Dims d = make_dims(N, S, M, D, L, Lq, P, padding);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = dtype == MSDA_F64 ? 8 : dtype == MSDA_F32 ? 4 : 2;
  const size_t nvalue = (size_t)N * S * M * D;
  if (nvalue == 0) flags &= ~MSDA_FLAG_DETERMINISTIC;   // nothing to accumulate: the plain kernels do
  const size_t ws_need = ws_main_bytes(N, S, M, D, dtype, flags);
  if (nvalue > 0 && (!grad_value || !value)) return MSDA_ERR_NULL_POINTER;
  if (ws_need > 0 && (!workspace || workspace_bytes < ws_need)) return MSDA_ERR_WORKSPACE;
  if (ws_need > 0 && !aligned16(workspace)) return MSDA_ERR_MISALIGNED;

  // CTA-tile kernels: the query order is the caller's plan (the workspace msda_forward_ex filled
  // for the same sampling locations) or is computed here into the tail of the workspace
  const bool fits32 = S * M * D < (int64_t(1) << 31);
  const bool aligned = aligned16(value) && aligned16(grad_output) && aligned16(grad_value) &&
                       aligned16(sampling_loc) && aligned16(attn_weight) &&
                       aligned16(grad_sampling_loc) && aligned16(grad_attn_weight);
  const bool fast = fits32 && fast_applies(D, L, dtype, flags) && aligned;
  const int* perm = nullptr;
  const char* saved = nullptr;   // the forward's saved tile plans behind the query order
  bool need_sort = false;
  if (N > 0 && Lq > 0 && fast && tile_applies(d, dtype, flags) && tile_encoder()) {
    if (plan && aligned16(plan)) {
      perm = (const int*)plan;
      if (save_applies(flags) && (reinterpret_cast<uintptr_t>(plan) & 127u) == 0)
        saved = (const char*)plan + tile_plan_bytes(N, Lq);
    } else if (workspace && aligned16(workspace) &&
               workspace_bytes >= round256(ws_need) + tile_plan_bytes(N, Lq)) {
      perm = (const int*)((char*)workspace + round256(ws_need));
      need_sort = true;
    }
  }

  if (flags & MSDA_FLAG_DETERMINISTIC) {
    // fixed-point image + header are always cleared here (MSDA_FLAG_SKIP_ZERO_FILL does not apply)
    cudaError_t e = cudaMemsetAsync(workspace, 0, ws_need, st);
    if (e != cudaSuccess) return cuda_fail(e);
    if (N == 0 || Lq == 0) {
      e = cudaMemsetAsync(grad_value, 0, nvalue * esz, st);
      return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
    }
    if (!spatial_shapes || !level_start_index || !sampling_loc || !attn_weight || !grad_output ||
        !grad_sampling_loc || !grad_attn_weight)
      return MSDA_ERR_NULL_POINTER;
    // the 8-lane vectorised instantiations cover the production head sizes; the 64-bit image
    // doubles the byte offsets the window kernels keep in 32 bits
    const bool vec8 = fast && ((dtype == MSDA_F32 && (D == 64 || D == 32)) || (dtype == MSDA_BF16 && D == 64));
    const int family = !vec8 ? 0
                       : (perm && Lq <= kSortStash) ? 3   // the order of a longer list is not reproducible
                       : (win_applies(d, dtype, flags) && S * M * D * 8 < (int64_t(1) << 32)) ? 2 : 1;
    if (family == 3 && need_sort) {
      const int r = launch_sort_dtype(dtype, sampling_loc, const_cast<int*>(perm), d, st);
      if (r != MSDA_OK) return r;
    }
#define MSDA_DET(T_, NU_)                                                                        \
  return backward_deterministic<T_, NU_>(value, spatial_shapes, level_start_index, sampling_loc,  \
                                         attn_weight, grad_output, grad_value, grad_sampling_loc, \
                                         grad_attn_weight, d, workspace, family, flags, st, perm, dtype)
    if (dtype == MSDA_F32 && D == 64) MSDA_DET(float, 2);
    if (dtype == MSDA_F32) MSDA_DET(float, 1);
    if (dtype == MSDA_F64) MSDA_DET(double, 1);
    MSDA_DET(__nv_bfloat16, 1);
#undef MSDA_DET
  }

  // zero the accumulation target (the reference does this with zeros_like,
  // cuda.cu:121-123; grad_loc / grad_aw are fully written by the kernels)
  void* acc = ws_need ? workspace : grad_value;
  if (nvalue > 0 && !(flags & MSDA_FLAG_SKIP_ZERO_FILL)) {
    const cudaError_t e = cudaMemsetAsync(acc, 0, ws_need ? ws_need : nvalue * esz, st);
    if (e != cudaSuccess) return cuda_fail(e);
  }
  if (N == 0 || Lq == 0) {
    // no query: grad_value is all zeros (for bf16 the image is not grad_value itself)
    if (nvalue > 0 && acc != grad_value) {
      const cudaError_t e = cudaMemsetAsync(grad_value, 0, nvalue * esz, st);
      if (e != cudaSuccess) return cuda_fail(e);
    }
    return MSDA_OK;
  }
  if (!spatial_shapes || !level_start_index || !sampling_loc || !attn_weight || !grad_output ||
      !grad_sampling_loc || !grad_attn_weight)
    return MSDA_ERR_NULL_POINTER;

  int r = MSDA_ERR_NO_TMA;
  if (perm) {
    if (need_sort) {
      r = launch_sort_dtype(dtype, sampling_loc, const_cast<int*>(perm), d, st);
      if (r != MSDA_OK) return r;
    }
    if (dtype == MSDA_BF16)
      // two 8-byte units per lane: dense 128-byte reds into the fp32 image (msda_fast.cuh)
      r = launch_bwd_tile<__nv_bfloat16, 2, float, UnitBf16x4>(
          value, spatial_shapes, level_start_index, sampling_loc, attn_weight, grad_output, perm,
          (float*)acc, grad_sampling_loc, grad_attn_weight, d, dtype, st, nullptr, need_sort,
          TileFront<__nv_bfloat16>(), saved);
    else if (D == 64)
      r = launch_bwd_tile<float, 2>(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                                    grad_output, perm, (float*)acc, grad_sampling_loc, grad_attn_weight,
                                    d, dtype, st, nullptr, need_sort, TileFront<float>(), saved);
    else
      r = launch_bwd_tile<float, 1>(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                                    grad_output, perm, (float*)acc, grad_sampling_loc, grad_attn_weight,
                                    d, dtype, st, nullptr, need_sort, TileFront<float>(), saved);
  }
  if (r != MSDA_ERR_NO_TMA) {
    // done by the tile kernels (or failed)
  } else if (fast && win_applies(d, dtype, flags)) {
    if (dtype == MSDA_BF16)
      // two 8-byte units per lane: dense 128-byte reds into the fp32 image (msda_fast.cuh)
      r = launch_bwd_win<__nv_bfloat16, 2, 8, float, UnitBf16x4>(
          value, spatial_shapes, level_start_index, sampling_loc, attn_weight, grad_output,
          (float*)acc, grad_sampling_loc, grad_attn_weight, d, st);
    else if (D == 64)
      r = launch_bwd_win<float, 2, 8>(value, spatial_shapes, level_start_index, sampling_loc,
                                      attn_weight, grad_output, (float*)acc, grad_sampling_loc,
                                      grad_attn_weight, d, st);
    else
      r = launch_bwd_win<float, 1, 8>(value, spatial_shapes, level_start_index, sampling_loc,
                                      attn_weight, grad_output, (float*)acc, grad_sampling_loc,
                                      grad_attn_weight, d, st);
  } else if (fast) {
    r = [&]() -> int {
      FastPlan plan;
      fast_plan(D, dtype, &plan);
      if (dtype == MSDA_BF16 && plan.nu == 1 && plan.g == 8)   // D = 64: dense reds, as above
        return launch_bwd_fast<__nv_bfloat16, 2, 8, float, UnitBf16x4>(
            value, spatial_shapes, level_start_index, sampling_loc, attn_weight, grad_output,
            (float*)acc, grad_sampling_loc, grad_attn_weight, d, flags, st);
      MSDA_DISPATCH_FAST(dtype, launch_bwd_fast, value, spatial_shapes, level_start_index,
                         sampling_loc, attn_weight, grad_output, (float*)acc, grad_sampling_loc,
                         grad_attn_weight, d, flags, st);
    }();
  } else if (dtype == MSDA_F32) {
    r = launch_bwd_generic<float, float>(value, spatial_shapes, level_start_index, sampling_loc,
                                         attn_weight, grad_output, (float*)acc, grad_sampling_loc,
                                         grad_attn_weight, d, st);
  } else if (dtype == MSDA_F64) {
    r = launch_bwd_generic<double, double>(value, spatial_shapes, level_start_index, sampling_loc,
                                           attn_weight, grad_output, (double*)acc,
                                           grad_sampling_loc, grad_attn_weight, d, st);
  } else {
    r = launch_bwd_generic<__nv_bfloat16, float>(value, spatial_shapes, level_start_index,
                                                 sampling_loc, attn_weight, grad_output,
                                                 (float*)acc, grad_sampling_loc, grad_attn_weight,
                                                 d, st);
  }
  if (r != MSDA_OK) return r;

  if (dtype == MSDA_BF16 && nvalue > 0) {
    if ((nvalue % 8) == 0 && aligned16(grad_value)) {
      const size_t n8 = nvalue / 8;
      size_t blocks = (n8 + 255) / 256;
      if (blocks > 148u * 16u) blocks = 148u * 16u;
      msda_round_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>((const float*)acc,
                                                               (__nv_bfloat16*)grad_value, n8);
    } else {
      size_t blocks = (nvalue + 255) / 256;
      if (blocks > 148u * 16u) blocks = 148u * 16u;
      msda_round_bf16_scalar_kernel<<<(unsigned)blocks, 256, 0, st>>>(
          (const float*)acc, (__nv_bfloat16*)grad_value, nvalue);
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e);
  }
  return MSDA_OK;
}

int msda_backward(const void* value, const int64_t* spatial_shapes, const int64_t* level_start_index,
                  const void* sampling_loc, const void* attn_weight, const void* grad_output,
                  void* grad_value, void* grad_sampling_loc, void* grad_attn_weight, int64_t N,
                  int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P, int dtype,
                  int padding, unsigned flags, void* workspace, size_t workspace_bytes,
                  void* stream) {
  return msda_backward_ex(value, spatial_shapes, level_start_index, sampling_loc, attn_weight,
                          grad_output, grad_value, grad_sampling_loc, grad_attn_weight, N, S, M, D, L,
                          Lq, P, dtype, padding, flags, workspace, workspace_bytes, nullptr, stream);
}


int msda_frontend_forward(const void* offsets, const void* logits, const void* ref,
                          const int64_t* level_sizes, void* loc, void* aw, int64_t N, int64_t Lq,
                          int64_t M, int64_t L, int64_t P, int ref_dim, int dtype, void* stream) {
  const int rc = check_front(N, Lq, M, L, P, ref_dim, dtype);
  if (rc != MSDA_OK) return rc;
  if (N == 0 || Lq == 0) return MSDA_OK;
  if (!offsets || !logits || !ref || !level_sizes || !loc || !aw) return MSDA_ERR_NULL_POINTER;
  FrontDims d;
  d.groups = (long long)(N * Lq * M);
  d.M = (int)M; d.L = (int)L; d.P = (int)P; d.ref_dim = ref_dim;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == MSDA_F32) return launch_front_fwd<float>(offsets, logits, ref, level_sizes, loc, aw, d, st);
  if (dtype == MSDA_F64) return launch_front_fwd<double>(offsets, logits, ref, level_sizes, loc, aw, d, st);
  return launch_front_fwd<__nv_bfloat16>(offsets, logits, ref, level_sizes, loc, aw, d, st);
}

int msda_frontend_backward(const void* offsets, const void* aw, const void* ref,
                           const int64_t* level_sizes, const void* grad_loc, const void* grad_aw,
                           void* grad_offsets, void* grad_logits, void* grad_ref, int64_t N,
                           int64_t Lq, int64_t M, int64_t L, int64_t P, int ref_dim, int dtype,
                           void* stream) {
  const int rc = check_front(N, Lq, M, L, P, ref_dim, dtype);
  if (rc != MSDA_OK) return rc;
  if (N == 0 || Lq == 0) return MSDA_OK;
  if (!offsets || !aw || !ref || !level_sizes || !grad_loc || !grad_aw || !grad_offsets || !grad_logits)
    return MSDA_ERR_NULL_POINTER;
  FrontDims d;
  d.groups = (long long)(N * Lq * M);
  d.M = (int)M; d.L = (int)L; d.P = (int)P; d.ref_dim = ref_dim;
  const int64_t ge = N * Lq * L * ref_dim;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == MSDA_F32)
    return launch_front_bwd<float>(offsets, aw, ref, level_sizes, grad_loc, grad_aw, grad_offsets,
                                   grad_logits, grad_ref, ge, d, st);
  if (dtype == MSDA_F64)
    return launch_front_bwd<double>(offsets, aw, ref, level_sizes, grad_loc, grad_aw, grad_offsets,
                                    grad_logits, grad_ref, ge, d, st);
  return launch_front_bwd<__nv_bfloat16>(offsets, aw, ref, level_sizes, grad_loc, grad_aw,
                                         grad_offsets, grad_logits, grad_ref, ge, d, st);
}


int msda_dam_scatter(const void* loc, const void* aw, const int64_t* level_sizes,
                     const int64_t* level_start, float* out, int64_t BL, int64_t Q, int64_t M,
                     int64_t L, int64_t P, int64_t S, int loc_cols, int dtype, void* stream) {
  if (BL < 0 || Q < 0 || M <= 0 || L <= 0 || P <= 0 || S < 0) return MSDA_ERR_BAD_SHAPE;
  if (loc_cols != 1 && loc_cols != 2) return MSDA_ERR_BAD_SHAPE;
  if (dtype != MSDA_F32 && dtype != MSDA_F64 && dtype != MSDA_BF16) return MSDA_ERR_BAD_DTYPE;
  if (S >= (int64_t(1) << 31) || M >= (int64_t(1) << 20) || L * P >= (int64_t(1) << 20) ||
      Q >= (int64_t(1) << 31))
    return MSDA_ERR_TOO_LARGE;
  if (BL > 0 && Q > 0 && BL * Q >= (int64_t(1) << 40) / (M * L * P)) return MSDA_ERR_TOO_LARGE;
  if (BL == 0 || S == 0) return MSDA_OK;
  if (!out) return MSDA_ERR_NULL_POINTER;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(out, 0, (size_t)(BL * M * S) * sizeof(float), st);
  if (e != cudaSuccess) return cuda_fail(e);
  if (Q == 0) return MSDA_OK;
  if (!loc || !aw || !level_sizes || !level_start) return MSDA_ERR_NULL_POINTER;
  DamDims d;
  d.points = (long long)(BL * Q * M * L * P);
  d.Q = (int)Q; d.M = (int)M; d.L = (int)L; d.P = (int)P; d.S = (int)S; d.loc_cols = loc_cols;
  const unsigned grid = (unsigned)((d.points + 255) / 256);
  if (dtype == MSDA_F32)
    msda_dam_scatter_kernel<float><<<grid, 256, 0, st>>>((const float*)loc, (const float*)aw,
                                                         level_sizes, level_start, out, d);
  else if (dtype == MSDA_F64)
    msda_dam_scatter_kernel<double><<<grid, 256, 0, st>>>((const double*)loc, (const double*)aw,
                                                          level_sizes, level_start, out, d);
  else
    msda_dam_scatter_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        (const __nv_bfloat16*)loc, (const __nv_bfloat16*)aw, level_sizes, level_start, out, d);
  e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}


}  // extern "C"

// ---- fused module entry points (second step) -----------
namespace {
template <typename T>
TileFront<T> make_front(const void* off, const void* logits, const void* ref, void* g_off, void* g_logits,
                        float* g_ref, int ref_dim, const unsigned char* mask) {
  TileFront<T> f;
  f.off = (const T*)off; f.logits = (const T*)logits; f.ref = (const T*)ref;
  f.g_off = (T*)g_off; f.g_logits = (T*)g_logits; f.g_ref = g_ref; f.ref_dim = ref_dim; f.mask = mask;
  return f;
}
inline bool module_supported(const Dims& d, int64_t S, int64_t M, int64_t D, int64_t L, int dtype, int ref_dim,
                             unsigned flags) {
  if (ref_dim != 1 && ref_dim != 2) return false;
  if (flags & (MSDA_FLAG_DETERMINISTIC | MSDA_FLAG_NO_TILE | MSDA_FLAG_NO_WINDOW | MSDA_FLAG_FORCE_GENERIC |
               MSDA_FLAG_PLAN))
    return false;
  if (dtype != MSDA_F32 && dtype != MSDA_BF16) return false;
  if (!(S * M * D < (int64_t(1) << 31)) || !fast_applies(D, L, dtype, flags)) return false;
  return tile_applies(d, dtype, flags) && tile_encoder() != nullptr;
}
}  // namespace

extern "C" {

int msda_module_supported(int64_t N, int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P,
                          int ref_dim, int dtype, unsigned flags) {
  if (check_common(N, S, M, D, L, Lq, P, dtype, MSDA_PAD_ZEROS) != MSDA_OK || N == 0 || Lq == 0) return 0;
  return module_supported(make_dims(N, S, M, D, L, Lq, P, MSDA_PAD_ZEROS), S, M, D, L, dtype, ref_dim, flags) ? 1 : 0;
}

int msda_module_forward(const void* value, const int64_t* spatial_shapes, const int64_t* level_start_index,
                        const void* offsets, const void* logits, const void* ref,
                        const unsigned char* padding_mask, void* output, int64_t N,
                        int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P, int ref_dim,
                        int dtype, int padding, unsigned flags, void* workspace, size_t workspace_bytes,
                        void* stream) {
  const int rc = check_common(N, S, M, D, L, Lq, P, dtype, padding);
  if (rc != MSDA_OK) return rc;
  if (N == 0 || Lq == 0) return MSDA_OK;
  if (!value || !spatial_shapes || !level_start_index || !offsets || !logits || !ref || !output)
    return MSDA_ERR_NULL_POINTER;
  Dims d = make_dims(N, S, M, D, L, Lq, P, padding);
  if (!module_supported(d, S, M, D, L, dtype, ref_dim, flags) || !workspace || !aligned16(workspace) ||
      workspace_bytes < tile_plan_bytes(N, Lq) || !aligned16(value) || !aligned16(output) || !aligned16(offsets) ||
      !aligned16(logits))
    return MSDA_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  int* perm = (int*)workspace;
  char* save = save_applies(flags) && workspace_bytes >= save_ws_bytes(N, M, Lq) &&
                       (reinterpret_cast<uintptr_t>(workspace) & 127u) == 0
                   ? (char*)workspace + tile_plan_bytes(N, Lq)
                   : nullptr;
  int r;
  if (dtype == MSDA_BF16) {
    const SortSrc<__nv_bfloat16> src = {nullptr, (const __nv_bfloat16*)offsets, (const __nv_bfloat16*)ref,
                                        spatial_shapes, ref_dim};
    if ((r = launch_sort<__nv_bfloat16>(src, perm, d, st)) != MSDA_OK) return r;
    r = launch_fwd_tile<__nv_bfloat16, 1, true>(value, spatial_shapes, level_start_index, nullptr, nullptr, perm,
                                                output, d, dtype, st, true,
                                                make_front<__nv_bfloat16>(offsets, logits, ref, nullptr, nullptr,
                                                                          nullptr, ref_dim, padding_mask),
                                                save);
  } else {
    const SortSrc<float> src = {nullptr, (const float*)offsets, (const float*)ref, spatial_shapes, ref_dim};
    if ((r = launch_sort<float>(src, perm, d, st)) != MSDA_OK) return r;
    const TileFront<float> f = make_front<float>(offsets, logits, ref, nullptr, nullptr, nullptr, ref_dim,
                                                 padding_mask);
    r = D == 64 ? launch_fwd_tile<float, 2, true>(value, spatial_shapes, level_start_index, nullptr, nullptr, perm,
                                                  output, d, dtype, st, true, f, save)
                : launch_fwd_tile<float, 1, true>(value, spatial_shapes, level_start_index, nullptr, nullptr, perm,
                                                  output, d, dtype, st, true, f, save);
  }
  return r == MSDA_ERR_NO_TMA ? MSDA_ERR_UNSUPPORTED : r;
}

int msda_module_backward(const void* value, const int64_t* spatial_shapes, const int64_t* level_start_index,
                         const void* offsets, const void* logits, const void* ref,
                         const unsigned char* padding_mask, const void* grad_output, void* grad_value, void* grad_offsets, void* grad_logits, float* grad_ref, int64_t N,
                         int64_t S, int64_t M, int64_t D, int64_t L, int64_t Lq, int64_t P, int ref_dim,
                         int dtype, int padding, unsigned flags, void* workspace, size_t workspace_bytes,
                         const void* plan, void* stream) {
  const int rc = check_common(N, S, M, D, L, Lq, P, dtype, padding);
  if (rc != MSDA_OK) return rc;
  if (N == 0 || Lq == 0) return MSDA_ERR_UNSUPPORTED;   // nothing fused about an empty call: use msda_backward
  if (!value || !spatial_shapes || !level_start_index || !offsets || !logits || !ref || !grad_output ||
      !grad_value || !grad_offsets || !grad_logits || !plan)
    return MSDA_ERR_NULL_POINTER;
  Dims d = make_dims(N, S, M, D, L, Lq, P, padding);
  const size_t ws_need = ws_main_bytes(N, S, M, D, dtype, flags);
  if (!module_supported(d, S, M, D, L, dtype, ref_dim, flags) || !aligned16(plan) || !aligned16(value) ||
      !aligned16(grad_output) || !aligned16(grad_value) || !aligned16(offsets) || !aligned16(logits) ||
      !aligned16(grad_offsets) || !aligned16(grad_logits))
    return MSDA_ERR_UNSUPPORTED;
  if (ws_need > 0 && (!workspace || workspace_bytes < ws_need)) return MSDA_ERR_WORKSPACE;
  if (ws_need > 0 && !aligned16(workspace)) return MSDA_ERR_MISALIGNED;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t esz = dtype == MSDA_F32 ? 4 : 2;
  const size_t nvalue = (size_t)N * S * M * D;
  void* acc = ws_need ? workspace : grad_value;
  cudaError_t e = cudaSuccess;
  if (!(flags & MSDA_FLAG_SKIP_ZERO_FILL)) e = cudaMemsetAsync(acc, 0, ws_need ? ws_need : nvalue * esz, st);
  if (e == cudaSuccess && grad_ref)
    e = cudaMemsetAsync(grad_ref, 0, (size_t)N * Lq * L * ref_dim * sizeof(float), st);
  if (e != cudaSuccess) return cuda_fail(e);
  const int* perm = (const int*)plan;
  const char* saved = save_applies(flags) && (reinterpret_cast<uintptr_t>(plan) & 127u) == 0
                          ? (const char*)plan + tile_plan_bytes(N, Lq)
                          : nullptr;
  int r;
  if (dtype == MSDA_BF16) {
    r = launch_bwd_tile<__nv_bfloat16, 2, float, UnitBf16x4, true>(
        value, spatial_shapes, level_start_index, nullptr, nullptr, grad_output, perm, (float*)acc, nullptr,
        nullptr, d, dtype, st, nullptr, false,
        make_front<__nv_bfloat16>(offsets, logits, ref, grad_offsets, grad_logits, grad_ref, ref_dim, padding_mask),
        saved);
  } else {
    const TileFront<float> f = make_front<float>(offsets, logits, ref, grad_offsets, grad_logits, grad_ref, ref_dim,
                                                 padding_mask);
    r = D == 64 ? launch_bwd_tile<float, 2, float, Unit<float>, true>(
                      value, spatial_shapes, level_start_index, nullptr, nullptr, grad_output, perm, (float*)acc,
                      nullptr, nullptr, d, dtype, st, nullptr, false, f, saved)
                : launch_bwd_tile<float, 1, float, Unit<float>, true>(
                      value, spatial_shapes, level_start_index, nullptr, nullptr, grad_output, perm, (float*)acc,
                      nullptr, nullptr, d, dtype, st, nullptr, false, f, saved);
  }
  if (r == MSDA_ERR_NO_TMA) return MSDA_ERR_UNSUPPORTED;
  if (r != MSDA_OK) return r;
  if (dtype == MSDA_BF16) {
    const size_t n8 = nvalue / 8;
    size_t blocks = (n8 + 255) / 256;
    if (blocks > 148u * 16u) blocks = 148u * 16u;
    msda_round_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>((const float*)acc, (__nv_bfloat16*)grad_value, n8);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e);
  }
  return MSDA_OK;
}

int msda_token_gather(const void* x, const int64_t* idx, const int64_t* nums, void* out, int64_t B,
                      int64_t S, int64_t K, int64_t row_bytes, void* stream) {
  if (B < 0 || S < 0 || K < 0 || row_bytes <= 0) return MSDA_ERR_BAD_SHAPE;
  if (B * K >= (int64_t(1) << 31) || B * S >= (int64_t(1) << 31) || row_bytes >= (int64_t(1) << 30))
    return MSDA_ERR_TOO_LARGE;
  if (B == 0 || K == 0) return MSDA_OK;
  if (!idx || !out || (S > 0 && !x)) return MSDA_ERR_NULL_POINTER;
  const unsigned grid = (unsigned)((B * K + 7) / 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (row_bytes % 16 == 0 && aligned16(x) && aligned16(out))
    msda_token_gather_kernel<16><<<grid, 256, 0, st>>>((const char*)x, idx, nums, (char*)out, (int)B, (int)S,
                                                       (int)K, (int)row_bytes);
  else
    msda_token_gather_kernel<1><<<grid, 256, 0, st>>>((const char*)x, idx, nums, (char*)out, (int)B, (int)S,
                                                      (int)K, (int)row_bytes);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

int msda_token_scatter(const void* base, const int64_t* idx, const int64_t* nums, const void* src, void* out,
                       int* inverse, int64_t B, int64_t S, int64_t K, int64_t row_bytes, int zero_hit_rows,
                       void* stream) {
  if (B < 0 || S < 0 || K < 0 || row_bytes <= 0) return MSDA_ERR_BAD_SHAPE;
  if (B * K >= (int64_t(1) << 31) || B * S >= (int64_t(1) << 31) || row_bytes >= (int64_t(1) << 30))
    return MSDA_ERR_TOO_LARGE;
  if (B == 0 || S == 0) return MSDA_OK;
  if (!base || !out || !inverse || (K > 0 && (!idx || (!src && !zero_hit_rows)))) return MSDA_ERR_NULL_POINTER;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(inverse, 0xff, (size_t)(B * S) * sizeof(int), st);   // -1: keep the base row
  if (e != cudaSuccess) return cuda_fail(e);
  if (K > 0)
    msda_token_inverse_kernel<<<(unsigned)((B * K + 255) / 256), 256, 0, st>>>(idx, nums, inverse, (int)B, (int)S,
                                                                               (int)K);
  const unsigned grid = (unsigned)((B * S + 7) / 8);
  const bool vec = row_bytes % 16 == 0 && aligned16(base) && aligned16(out) && (zero_hit_rows || aligned16(src));
#define MSDA_SCATTER(V_, Z_)                                                                               \
  msda_token_scatter_kernel<V_, Z_><<<grid, 256, 0, st>>>((const char*)base, inverse, (const char*)src,  \
                                                          (char*)out, (int)B, (int)S, (int)K, (int)row_bytes)
  if (vec && zero_hit_rows) MSDA_SCATTER(16, true);
  else if (vec) MSDA_SCATTER(16, false);
  else if (zero_hit_rows) MSDA_SCATTER(1, true);
  else MSDA_SCATTER(1, false);
#undef MSDA_SCATTER
  e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}


int msda_prepare_encoder_inputs(const void* const* srcs, const void* const* pos_embeds,
                                const unsigned char* const* masks, const int64_t* level_sizes_host,
                                const void* level_embed, void* src_flatten, void* lvl_pos_embed_flatten,
                                unsigned char* mask_flatten, float* valid_ratios, int64_t B, int64_t C,
                                int64_t L, int dtype, void* stream) {
  if (B < 0 || C <= 0 || L <= 0 || L > kPrepMaxL) return MSDA_ERR_BAD_SHAPE;
  if (dtype != MSDA_F32 && dtype != MSDA_F64 && dtype != MSDA_BF16) return MSDA_ERR_BAD_DTYPE;
  if (!srcs || !pos_embeds || !masks || !level_sizes_host) return MSDA_ERR_NULL_POINTER;
  PrepLevels lv;
  lv.L = (int)L;
  int64_t S = 0;
  int tiles = 0;
  for (int l = 0; l < (int)L; ++l) {
    const int64_t Tl = level_sizes_host[l];
    if (Tl < 0 || S + Tl >= (int64_t(1) << 31)) return MSDA_ERR_TOO_LARGE;
    if (Tl > 0 && B > 0 && (!srcs[l] || !pos_embeds[l] || !masks[l])) return MSDA_ERR_NULL_POINTER;
    lv.src[l] = srcs[l]; lv.pos[l] = pos_embeds[l]; lv.mask[l] = masks[l];
    lv.T[l] = (int)Tl; lv.start[l] = (int)S; lv.tile0[l] = tiles;
    S += Tl;
    tiles += (int)((Tl + 31) / 32);
  }
  lv.tile0[L] = tiles;
  if (B == 0 || S == 0) return MSDA_OK;
  if (!level_embed || !src_flatten || !lvl_pos_embed_flatten || !mask_flatten || !valid_ratios)
    return MSDA_ERR_NULL_POINTER;
  if (B > 65535 || (C + 31) / 32 > 65535) return MSDA_ERR_TOO_LARGE;
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid((unsigned)tiles, (unsigned)((C + 31) / 32), (unsigned)B);
#define MSDA_PREP(T_)                                                                                   \
  msda_prepare_encoder_kernel<T_><<<grid, 256, 0, st>>>(lv, (const T_*)level_embed, (T_*)src_flatten, \
                                                        (T_*)lvl_pos_embed_flatten, mask_flatten, (int)B, (int)C, (int)S)
  if (dtype == MSDA_F32) MSDA_PREP(float);
  else if (dtype == MSDA_F64) MSDA_PREP(double);
  else MSDA_PREP(__nv_bfloat16);
#undef MSDA_PREP
  msda_valid_ratio_kernel<<<dim3((unsigned)L, (unsigned)B), 32, 0, st>>>(lv, valid_ratios, (int)B);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

int msda_encoder_reference_points(const int64_t* level_sizes, const float* valid_ratios, float* out,
                                  int64_t N, int64_t S, int64_t L, void* stream) {
  if (N < 0 || S < 0 || L <= 0) return MSDA_ERR_BAD_SHAPE;
  if (S >= (int64_t(1) << 31) || L >= (int64_t(1) << 20) || (N > 0 && S > 0 && N >= (int64_t(1) << 40) / S))
    return MSDA_ERR_TOO_LARGE;
  if (N == 0 || S == 0) return MSDA_OK;
  if (!level_sizes || !valid_ratios || !out) return MSDA_ERR_NULL_POINTER;
  const long long total = (long long)(N * S);
  msda_reference_points_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      level_sizes, valid_ratios, out, total, (int)S, (int)L);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MSDA_OK : cuda_fail(e);
}

}  // extern "C"
