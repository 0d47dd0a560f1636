// msda_plan.cuh -- the CTA-tile path split in two: a PLAN kernel and APPLY kernels.
//
// Why: in the one-kernel form (msda_tile.cuh) a CTA spends 60 % of its life in the prologue
// (points from HBM, geometry, window agreement, weight scatter) at 3-4 warps per scheduler,
// every instruction waiting ~9 cycles on the one before; the row loops, which are the actual
// gather, get 25 % (profiles/r2_*).  More warps do not fit next to a 40 KB value tile and 128
// registers.  Split:
//   msda_plan_kernel   no value tile, <= 64 registers, 32 warps per SM: everything the prologue
//                      computed goes to a PLAN in global memory (13 KB per tile of 64 queries:
//                      the tile's row windows, the query list, per block its windows and the
//                      merged weight matrix Wt).  Written once, read by the forward AND the
//                      backward of a training step (it stays in the 126 MB L2 in between).
//   msda_*_apply_kernel  one thread turns the tile header into TMA boxes for the value rows and
//                      bulk copies for the weights; the warps go straight to the row loops.
// Tile geometry, block layout and the row loops are those of msda_tile.cuh.
#pragma once
#include "msda_tile.cuh"

namespace msda {

// ---- plan layout (per tile, kPlanStride bytes) --------------------------------------
//   [0, 64)      int tlo[4], tbase[4], total, mode, rtw[kTileWarps] (packed rows per warp), pad
//   [64, 64+4Q)  int q[kTileQ]          the tile's queries in time order (-1: none)
//   [.., +32B)   per block: int blo[4]; u8 brows[4]; u8 wmax[4]; u8 wrows[4]; pad
//   [kPlanMetaBytes, ..)  Wt: per warp a region of 4 * kTileBlkRows rows of 16 B; block k of the
//                warp at row k * rtw (compact, so that one bulk copy moves a warp's weights)
constexpr int kPlanModeTile = 1;      // plan valid: row loops
constexpr int kPlanModeTap2 = 0;      // tile does not fit: per-tap body, temporal maps
constexpr int kPlanModeTap4 = 2;      // maps with H != 1: per-tap body, 4 taps
constexpr int kPlanHdrBytes = 64;
constexpr int kPlanQOff = kPlanHdrBytes;
constexpr int kPlanBlkOff = kPlanQOff + 4 * kTileQ;
constexpr int kPlanBlkBytes = 32;
constexpr int kPlanMetaBytes = (kPlanBlkOff + kTileBlocks * kPlanBlkBytes + 127) / 128 * 128;
constexpr int kPlanWarpWtBytes = 4 * kTileBlkRows * 16;
constexpr int kPlanWtBytes = kTileWarps * kPlanWarpWtBytes;
constexpr int kPlanStride = kPlanMetaBytes + kPlanWtBytes;
static_assert(kTileWarps <= 6, "header holds rtw[] behind 10 ints");

struct __align__(16) PlanBlk {
  int blo[4];
  unsigned char brows[4], wmax[4], wrows[4], pad[4];
};
static_assert(sizeof(PlanBlk) == kPlanBlkBytes, "plan layout");

// ------------------------------- plan ----------------------------------------
// One sampling point as the plan kernel sees it: pixel coordinate (clamped in border mode) and
// the attention weight times the vertical factor (0 when the point is gated off).  The weight
// the point gives slab row r of its level is  aw' * max(0, 1 - |w_im - (r - start)|)  -- the two
// taps of msda_win.cuh's win_point() are the two rows where that hat is non-zero; rows outside
// the map are never asked for, which is the zero padding of the reference (.cuh:56-79).
__device__ __forceinline__ void plan_point(float x, float y, float a, int W, bool border, bool active,
                                           float& w_im, float& awf, int& r_lo, bool& live) {
  float h_im = y - 0.5f;
  w_im = x * (float)W - 0.5f;
  if (border) {
    w_im = fminf(fmaxf(w_im, 0.f), (float)(W - 1));   // a NaN clamps to column 0
    h_im = 0.f;
    live = active;
  } else {
    live = active && (fabsf(h_im) < 1.f) && (w_im > -1.f) && (w_im < (float)W);
  }
  awf = live ? a * (1.f - fabsf(h_im)) : 0.f;
  r_lo = (int)floorf(w_im);
}

// No shared-memory weight matrix: lane (qi, h) evaluates column (level, qi) of its block's Wt row
// by row and stores the floats where the apply kernels' bulk copies expect them.
template <typename T>
__global__ void __launch_bounds__(kTileThreads, 8)
msda_plan_kernel(const int64_t* __restrict__ shapes, const int64_t* __restrict__ lsi,
                 const T* __restrict__ loc, const T* __restrict__ aw, const int* __restrict__ perm,
                 char* __restrict__ plan, Dims d, int nchunks) {
  constexpr unsigned FULL = 0xffffffffu;
  __shared__ int s_lvl[3 * kWinMaxL];
  __shared__ int s_win[2 * kWinMaxL];
  __shared__ int s_q[kTileQ];
  __shared__ __align__(8) unsigned long long s_mbar;   // unused here (tile_setup initialises it)
  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7, warp = tid >> 5, k = g & 3;
  const int qi = j & 3, h = j >> 2, gbase = (tid & 31) & ~7;
  const int m = blockIdx.x % d.M;
  const int t2 = blockIdx.x / d.M;
  const int chunk = t2 % nchunks, b = t2 / nchunks;
  char* tp = plan + (size_t)blockIdx.x * kPlanStride;
  int* hdr = reinterpret_cast<int*>(tp);
  const bool one_d = tile_setup(shapes, lsi, perm, d, b, chunk, s_lvl, s_win, s_q, &s_mbar);
  // the apply kernels of this call may be scheduled now; they wait for this grid's completion
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (tid < kTileQ) reinterpret_cast<int*>(tp + kPlanQOff)[tid] = s_q[tid];
  if (!one_d) {
    if (tid == 0) hdr[9] = kPlanModeTap4;
    return;
  }
  const int q = s_q[g * kWinQ + qi];
  const bool have_q = q >= 0;
  const int LP = d.L * kWinP;
  const size_t grp = ((size_t)b * d.Lq + (have_q ? q : 0)) * d.M + m;

  float wim[2][4], awf[2][4];
  int lo[2], hi[2], start[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int l = h + 2 * i;
    const bool active = have_q && l < d.L;
    const int ls = l < d.L ? l : 0;
    float x[4], y[4], a[4];
    WinPts<T>::ld(loc + (grp * LP + ls * kWinP) * 2, aw + grp * LP + ls * kWinP, x, y, a);
    const int W = s_lvl[3 * ls + 1];
    start[i] = s_lvl[3 * ls + 2];
    lo[i] = kTileBig; hi[i] = -kTileBig;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      int r;
      bool live;
      plan_point(x[p], y[p], a[p], W, d.border != 0, active, wim[i][p], awf[i][p], r, live);
      lo[i] = min(lo[i], live ? r : kTileBig);
      hi[i] = max(hi[i], live ? r + 1 : -kTileBig);
    }
    // rows relative to the level, clipped to the map; then slab rows
    lo[i] = max(lo[i], 0) + start[i];
    hi[i] = min(hi[i], W - 1) + start[i];
    if (hi[i] < lo[i]) { lo[i] = kTileBig; hi[i] = -kTileBig; }
  }
#pragma unroll
  for (int s = 1; s <= 2; s <<= 1)
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      lo[i] = min(lo[i], __shfl_xor_sync(FULL, lo[i], s));
      hi[i] = max(hi[i], __shfl_xor_sync(FULL, hi[i], s));
    }
  if (qi == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i)
      if (hi[i] >= lo[i]) {
        atomicMin(&s_win[h + 2 * i], lo[i]);
        atomicMax(&s_win[kWinMaxL + h + 2 * i], hi[i]);
      }
  }
  int blo[kWinMaxL], brows[kWinMaxL], wmax[kWinMaxL], wrows[kWinMaxL], bpoff[kWinMaxL];
  int rt = 0;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int src = gbase + (l & 1) * 4;
    const int llo = __shfl_sync(FULL, lo[l >> 1], src);
    const int lhi = __shfl_sync(FULL, hi[l >> 1], src);
    const int rows = max(lhi - llo + 1, 0);
    blo[l] = llo;
    brows[l] = rows;
    wmax[l] = __reduce_max_sync(FULL, rows);
    wrows[l] = (wmax[l] + 1) & ~1;
    bpoff[l] = rt;
    rt += wrows[l];
  }
  const bool over = __syncthreads_or(rt > kTileBlkRows) != 0;

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
  const bool ok = !over && total <= kTileCapRows;
  if (tid == 0) {
#pragma unroll
    for (int l = 0; l < kWinMaxL; ++l) { hdr[l] = tlo[l]; hdr[4 + l] = tbase[l]; }
    hdr[8] = total;
    hdr[9] = ok ? kPlanModeTile : kPlanModeTap2;
  }
  if (!ok) return;
  if ((tid & 31) == 0) hdr[10 + warp] = rt;
  if (j == 0) {
    PlanBlk pb;
#pragma unroll
    for (int l = 0; l < kWinMaxL; ++l) {
      pb.blo[l] = blo[l];
      pb.brows[l] = (unsigned char)brows[l];
      pb.wmax[l] = (unsigned char)wmax[l];
      pb.wrows[l] = (unsigned char)wrows[l];
      pb.pad[l] = (unsigned char)(brows[l] > 0 ? tbase[l] + blo[l] - tlo[l] : 0);   // row of the value tile
    }
    *reinterpret_cast<PlanBlk*>(tp + kPlanBlkOff + g * kPlanBlkBytes) = pb;
  }
  // weights: compact per warp (block k at row k * rt); lanes qi = 0..3 fill one 16-byte row
  float* wg = reinterpret_cast<float*>(tp + kPlanMetaBytes + warp * kPlanWarpWtBytes) + (k * rt) * 4 + qi;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int my_lo = h ? blo[2 * i + 1] : blo[2 * i];
    const int my_rows = h ? brows[2 * i + 1] : brows[2 * i];
    const int my_n = h ? wrows[2 * i + 1] : wrows[2 * i];
    const int my_off = h ? bpoff[2 * i + 1] : bpoff[2 * i];
    const float r_first = (float)(my_lo - start[i]);
    for (int t = 0; t < my_n; ++t) {
      const float r = r_first + (float)t;
      float w = 0.f;
#pragma unroll
      for (int p = 0; p < 4; ++p) w = fmaf(awf[i][p], fmaxf(1.f - fabsf(wim[i][p] - r), 0.f), w);
      wg[(my_off + t) * 4] = t < my_rows ? w : 0.f;
    }
  }
}

// ------------------------------- apply: common ---------------------------------
// Persistent CTAs (two per SM), each a pipeline of kApplyStages shared-memory stages:
//   producer warp (one thread): tile header from the plan -> one mbarrier transaction per stage:
//       bulk copy of the tile's meta block (header, query list, block descriptors), TMA boxes of
//       the value rows, bulk copies of the warps' weights;
//   kTileWarps consumer warps: wait for the stage, run their blocks' row loops out of shared
//       memory, write their results, release the stage.
// No CTA barrier in the loop; the value rows of tile i+1 fly in while tile i is computed.
constexpr int kApplyStages = 2;
constexpr int kApplyThreads = kTileThreads + 32;

template <int ROW_BYTES>
struct ApplySmem {
  static constexpr int kValueBytes = kTileCapRows * ROW_BYTES;
  static constexpr int kStageBytes = kValueBytes + kPlanWtBytes + kPlanMetaBytes;
  static constexpr int kBytes = kApplyStages * kStageBytes;
};
static_assert(kTileCapRows <= 255, "tile rows are stored in a byte");
static_assert(kPlanMetaBytes % 128 == 0 && kPlanWtBytes % 128 == 0, "stage sections stay 128-byte aligned");

__device__ __forceinline__ void tile_mbar_arrive(unsigned long long* mbar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tile_smem_u32(mbar)) : "memory");
}

// L2 prefetch of one value box (no shared memory involved)
__device__ __forceinline__ void tile_tma_prefetch(const CUtensorMap* map, int col, int row) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(
                   reinterpret_cast<unsigned long long>(map)),
               "r"(col), "r"(row)
               : "memory");
}

struct ApplyHdr {
  int4 h0, h1, h2, h3;   // tlo | tbase | total, mode, rtw[0], rtw[1] | rtw[2..]
};
__device__ __forceinline__ void apply_load_hdr(const char* plan, int tile, int ntiles, ApplyHdr& h) {
  if (tile < ntiles) {
    const char* tp = plan + (size_t)tile * kPlanStride;
    h.h0 = tile_ld_dep4(tp);
    h.h1 = tile_ld_dep4(tp + 16);
    h.h2 = tile_ld_dep4(tp + 32);
    h.h3 = tile_ld_dep4(tp + 48);
  }
}
// the value boxes of a tile: into the stage (PREFETCH = false) or into L2 only
template <bool PREFETCH>
__device__ __forceinline__ void apply_value_boxes(const CUtensorMap* vmap, const ApplyHdr& h, const Dims& d,
                                                  int nchunks, int tile, char* st, unsigned row_bytes,
                                                  unsigned long long* mbar) {
  const int total = h.h2.x;
  const int m = tile % d.M, b = tile / d.M / nchunks;
  const int tlo[4] = {h.h0.x, h.h0.y, h.h0.z, h.h0.w}, tbase[4] = {h.h1.x, h.h1.y, h.h1.z, h.h1.w};
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int nb = ((l + 1 < kWinMaxL ? tbase[l + 1] : total) - tbase[l]) / kTileBoxRows;
    for (int k = 0; k < nb; ++k) {
      if (PREFETCH)
        tile_tma_prefetch(vmap, m * d.D, b * d.S + tlo[l] + k * kTileBoxRows);
      else
        tile_tma_load(st + (size_t)(tbase[l] + k * kTileBoxRows) * row_bytes, vmap, m * d.D,
                      b * d.S + tlo[l] + k * kTileBoxRows, mbar);
    }
  }
}

// The producer: fills stage after stage.  Called by ONE thread.  Headers are read two tiles
// ahead and the value rows of the next tile are pulled into L2 while the current one is loaded,
// so that the load into a stage that has just been released is an L2 hit.
__device__ __forceinline__ void apply_producer(const CUtensorMap* vmap, const char* plan, const Dims& d,
                                               int nchunks, int ntiles, char* stages, int stage_bytes,
                                               int value_bytes, unsigned row_bytes,
                                               unsigned long long* full, unsigned long long* empty) {
  ApplyHdr cur, nxt;
  apply_load_hdr(plan, blockIdx.x, ntiles, cur);
  apply_load_hdr(plan, blockIdx.x + gridDim.x, ntiles, nxt);
  int it = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const int s = it % kApplyStages, use = it / kApplyStages;
    const int tile_n = tile + gridDim.x;
    if (tile_n < ntiles && nxt.h2.y == kPlanModeTile && nxt.h2.x > 0)
      apply_value_boxes<true>(vmap, nxt, d, nchunks, tile_n, nullptr, row_bytes, nullptr);
    if (use > 0) tile_mbar_wait(&empty[s], (use - 1) & 1);
    const char* tp = plan + (size_t)tile * kPlanStride;
    char* st = stages + s * stage_bytes;
    const int total = cur.h2.x, mode = cur.h2.y;
    const int rtw[6] = {cur.h2.z, cur.h2.w, cur.h3.x, cur.h3.y, cur.h3.z, cur.h3.w};
    unsigned bytes = kPlanMetaBytes;
    const bool rows = mode == kPlanModeTile && total > 0;
    if (rows) {
      bytes += (unsigned)total * row_bytes;
#pragma unroll
      for (int w = 0; w < kTileWarps; ++w) bytes += (unsigned)rtw[w] * 64u;
    }
    tile_mbar_expect_tx(&full[s], bytes);
    tile_bulk_load(st + value_bytes + kPlanWtBytes, tp, kPlanMetaBytes, &full[s]);
    if (rows) {
      apply_value_boxes<false>(vmap, cur, d, nchunks, tile, st, row_bytes, &full[s]);
#pragma unroll
      for (int w = 0; w < kTileWarps; ++w)
        if (rtw[w] > 0)
          tile_bulk_load(st + value_bytes + w * kPlanWarpWtBytes, tp + kPlanMetaBytes + w * kPlanWarpWtBytes,
                         (unsigned)rtw[w] * 64u, &full[s]);
    }
    cur = nxt;
    apply_load_hdr(plan, tile_n + gridDim.x, ntiles, nxt);
  }
}

// per-thread block description, read from the stage's meta block
struct ApplyBlk {
  int blo[kWinMaxL], brows[kWinMaxL], wmax[kWinMaxL], wrows[kWinMaxL], trow[kWinMaxL], bpoff[kWinMaxL];
  int rt;
};
__device__ __forceinline__ void apply_block(const char* meta, int g, int warp, ApplyBlk& ab) {
  const int4 b0 = *reinterpret_cast<const int4*>(meta + kPlanBlkOff + g * kPlanBlkBytes);
  const uint4 b1 = *reinterpret_cast<const uint4*>(meta + kPlanBlkOff + g * kPlanBlkBytes + 16);
  const int blo[4] = {b0.x, b0.y, b0.z, b0.w};
  int off = 0;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    ab.blo[l] = blo[l];
    ab.brows[l] = (b1.x >> (8 * l)) & 255;
    ab.wmax[l] = (b1.y >> (8 * l)) & 255;
    ab.wrows[l] = (b1.z >> (8 * l)) & 255;
    ab.trow[l] = (b1.w >> (8 * l)) & 255;
    ab.bpoff[l] = off;
    off += ab.wrows[l];
  }
  ab.rt = reinterpret_cast<const int*>(meta)[10 + warp];
}

// ------------------------------- forward apply ---------------------------------
template <typename T, int NU>
__global__ void __launch_bounds__(kApplyThreads, 2)
msda_fwd_apply_kernel(const __grid_constant__ CUtensorMap vmap, const T* __restrict__ value,
                      const int64_t* __restrict__ shapes, const int64_t* __restrict__ lsi,
                      const T* __restrict__ loc, const T* __restrict__ aw,
                      const char* __restrict__ plan, T* __restrict__ out, Dims d, int nchunks,
                      int ntiles) {
  using U = Unit<T>;
  using Raw = typename U::Raw;
  constexpr int G = 8;
  constexpr int E = U::E;
  constexpr int RB = NU * G * 16;
  using SM = ApplySmem<RB>;
  __shared__ int s_lvl[3 * kWinMaxL];
  __shared__ __align__(8) unsigned long long s_full[kApplyStages], s_empty[kApplyStages];
  extern __shared__ __align__(128) char s_raw[];
  const int tid = threadIdx.x, g = tid >> 3, j = tid & 7, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kApplyStages; ++s) {
      tile_mbar_init(&s_full[s], 1);
      tile_mbar_init(&s_empty[s], kTileWarps);
    }
  }
  win_stage_levels(shapes, lsi, d.L, s_lvl);   // also the barrier behind the mbarrier inits
  tile_grid_wait();
  if (warp == kTileWarps) {
    if ((tid & 31) == 0)
      apply_producer(&vmap, plan, d, nchunks, ntiles, s_raw, SM::kStageBytes, SM::kValueBytes, RB, s_full,
                     s_empty);
    return;
  }
  int it = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const int s = it % kApplyStages, use = it / kApplyStages;
    char* vtile = s_raw + s * SM::kStageBytes;
    char* s_wt = vtile + SM::kValueBytes;
    const char* meta = s_wt + kPlanWtBytes;
    const int* s_q = reinterpret_cast<const int*>(meta + kPlanQOff);
    const int m = tile % d.M, b = tile / d.M / nchunks;
    tile_mbar_wait(&s_full[s], use & 1);
    const int mode = reinterpret_cast<const int*>(meta)[9];
    if (mode != kPlanModeTile) {
      if (mode == kPlanModeTap4) tile_fallback_fwd<T, NU, 4>(value, loc, aw, out, d, b, m, s_lvl, s_wt, s_q);
      else tile_fallback_fwd<T, NU, 2>(value, loc, aw, out, d, b, m, s_lvl, s_wt, s_q);
    } else {
      ApplyBlk ab;
      apply_block(meta, g, warp, ab);
      float2 acc[kWinQ][NU][E / 2];
#pragma unroll
      for (int q = 0; q < kWinQ; ++q)
#pragma unroll
        for (int uu = 0; uu < NU; ++uu)
#pragma unroll
          for (int c = 0; c < E / 2; ++c) acc[q][uu][c] = make_float2(0.f, 0.f);
      const float4* Wt = reinterpret_cast<const float4*>(s_wt + warp * kPlanWarpWtBytes) + (g & 3) * ab.rt;
#pragma unroll
      for (int l = 0; l < kWinMaxL; ++l) {
        const char* vr = vtile + ab.trow[l] * RB + j * 16;
        const float4* wl = Wt + ab.bpoff[l];
        const int n = ab.wmax[l], nb = ab.brows[l];
#pragma unroll 2
        for (int i = 0; i < n; ++i) {
          const float4 wv = wl[i];
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
    // this warp is done with the stage (the per-tap body used it as scratch through the generic
    // proxy: order that before the producer's next asynchronous copies)
    if (mode != kPlanModeTile) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if ((tid & 31) == 0) tile_mbar_arrive(&s_empty[s]);
  }
}

}  // namespace msda
