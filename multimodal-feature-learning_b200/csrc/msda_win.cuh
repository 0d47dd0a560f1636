// msda_win.cuh -- "row window" MSDeformAttn kernels for the temporal (H == 1)
// use of the op when consecutive queries sample neighbouring rows (encoder
// self-attention: models/deformable/unimodal_deformable_transformer.py:259-276
// builds reference points that are monotone in the token index, and the
// module's offsets are a few rows, models/ops/modules/ms_deform_attn.py:64-71).
//
// A BLOCK = kWinQ = 4 consecutive queries of one (batch, head), owned by a
// group of G lanes; lane j owns NU*E channels of the head row.
//   1. geometry: lane (qi = j & 3, h = j >> 2) owns the 4 points of query qi
//      at the levels l = h, h + G/4, ...; per level the block agrees on the
//      window [lo_l, hi_l] of rows any of its taps touches (shuffles).
//   2. the windows of the (<= 4) levels are packed back to back (Rtot rows) and
//      the interpolation * attention weights are scattered into a small
//      shared-memory matrix Wt[t][qi] (column (level, qi) is private to one
//      lane: plain read-modify-write, no atomics), next to rowoff[t] = element
//      offset of the value row behind packed row t.
//   3. forward : out[qi] += Wt[t][qi] * value[row t] for t < Rtot -- each value
//                row is loaded ONCE per block (not once per tap: 7.7 instead
//                of 32 row loads per query at the encoder shape, and, because
//                taps that share a row merge, no more FMAs than per-tap code);
//      backward: per row t, in registers,
//                  grad_value[row t] += sum_qi Wt[t][qi] * grad_out[qi]   (ONE 128-bit
//                                        red per row and lane, not one per tap)
//                  G[t][qi]           = <grad_out[qi], value[row t]>     (transpose-
//                                        reduce over the G lanes, stored over Wt)
//                and afterwards every point picks its two dots G[t0][qi],
//                G[t0+1][qi] to form grad_attn_weight / grad_sampling_loc
//                (.cuh:115-159 for the math).
// A block whose packed windows exceed kWinCap rows (random locations, a block
// straddling two query levels) runs the per-tap body of msda_fast.cuh on its
// 4 queries instead; maps with H != 1 always do.
//
// Note: with the row-window form a non-finite value row poisons the other
// queries of the same 4-block through 0 * inf (the per-tap kernels never
// read a row they do not tap).  MSDA_FLAG_NO_WINDOW selects the per-tap kernels.
#pragma once
#include "msda_fast.cuh"

namespace msda {

constexpr int kWinQ = 4;         // queries per block
constexpr int kWinP = 4;         // points per level (template-free fast case)
constexpr int kWinMaxL = 4;      // levels
constexpr int kWinCap = 72;      // packed rows per block, multiple of 8 (the encoder shape peaks at 70 rows)
constexpr int kWinFillPad = 8;   // the row loops run to a multiple of this and read that far ahead
constexpr int kWinRows = kWinCap + kWinFillPad;  // + padding rows read by the software pipeline
constexpr int kWinThreads = 128;     // 4 warps = 4 adjacent units per CTA (they share rows in L1)
constexpr int kWinMinQueries = 256;  // host heuristic: shorter query lists use the per-tap kernels

struct WinCoef { float k0, k1, fx, cy0, cy1; };

// One sampling point on a (1, W) map: slab row of the low tap and the tap
// weights (already times the attention weight and the vertical factor; 0 for a
// tap that is not read).  Same arithmetic as point_geom() with H == 1
// (msda_common.cuh): h_im = y - 0.5 in (-1, 1) -> the single row is the low
// (weight 1 - h_im) or the high (weight h_im + 1) row of the vertical pair,
// i.e. 1 - |h_im| either way.
template <bool BWD>
__device__ __forceinline__ void win_point(float x, float y, float a, int W, int start, bool border,
                                          bool active, int& r0, float& w0, float& w1, bool& live,
                                          bool& v0, bool& v1, WinCoef& c) {
  float h_im = y - 0.5f;
  float w_im = x * (float)W - 0.5f;
  const float wlast = (float)(W - 1);
  float sx = (float)W, sy = 1.f;
  if (border) {
    if (!(w_im > 0.f)) { w_im = 0.f; sx = 0.f; }
    else if (w_im >= wlast) { w_im = wlast; sx = 0.f; }
    h_im = 0.f; sy = 0.f;
    live = active;
  } else {
    live = active && (fabsf(h_im) < 1.f) && (w_im > -1.f) && (w_im < (float)W);
  }
  const float wf = floorf(w_im);
  const float lw = w_im - wf, hw = 1.f - lw;
  const float fy = 1.f - fabsf(h_im);
  v0 = live && w_im >= 0.f;      // w_low >= 0
  v1 = live && w_im < wlast;     // w_low + 1 <= W - 1
  r0 = start + (int)wf;
  const float k0 = fy * hw, k1 = fy * lw;
  w0 = v0 ? k0 * a : 0.f;
  w1 = v1 ? k1 * a : 0.f;
  if (BWD) {
    c.k0 = v0 ? k0 : 0.f;
    c.k1 = v1 ? k1 : 0.f;
    c.fx = live ? sx * a * fy : 0.f;
    const float sg = copysignf(sy, -h_im) * a;   // h_im >= 0: -sy, else +sy
    c.cy0 = live ? sg * hw : 0.f;                // (a NaN location has NaN hw / lw)
    c.cy1 = live ? sg * lw : 0.f;
  }
}

// (x, y, attention weight) of the 4 points of one (query, level)
template <typename T> struct WinPts;
template <> struct WinPts<float> {
  static __device__ __forceinline__ void ld(const float* loc2, const float* awp, float (&x)[4],
                                            float (&y)[4], float (&a)[4]) {
    const float4 l0 = __ldg(reinterpret_cast<const float4*>(loc2));
    const float4 l1 = __ldg(reinterpret_cast<const float4*>(loc2) + 1);
    const float4 w = __ldg(reinterpret_cast<const float4*>(awp));
    x[0] = l0.x; y[0] = l0.y; x[1] = l0.z; y[1] = l0.w;
    x[2] = l1.x; y[2] = l1.y; x[3] = l1.z; y[3] = l1.w;
    a[0] = w.x; a[1] = w.y; a[2] = w.z; a[3] = w.w;
  }
  // read-once variant: no L1 allocation
  static __device__ __forceinline__ void ld_stream(const float* loc2, const float* awp, float (&x)[4],
                                                   float (&y)[4], float (&a)[4]) {
    const float4 l0 = __ldcs(reinterpret_cast<const float4*>(loc2));
    const float4 l1 = __ldcs(reinterpret_cast<const float4*>(loc2) + 1);
    const float4 w = __ldcs(reinterpret_cast<const float4*>(awp));
    x[0] = l0.x; y[0] = l0.y; x[1] = l0.z; y[1] = l0.w;
    x[2] = l1.x; y[2] = l1.y; x[3] = l1.z; y[3] = l1.w;
    a[0] = w.x; a[1] = w.y; a[2] = w.z; a[3] = w.w;
  }
  static __device__ __forceinline__ void ld_smem(const float* loc2, const float* awp, float (&x)[4],
                                                 float (&y)[4], float (&a)[4]) {
    const float4 l0 = *reinterpret_cast<const float4*>(loc2);
    const float4 l1 = *(reinterpret_cast<const float4*>(loc2) + 1);
    const float4 w = *reinterpret_cast<const float4*>(awp);
    x[0] = l0.x; y[0] = l0.y; x[1] = l0.z; y[1] = l0.w;
    x[2] = l1.x; y[2] = l1.y; x[3] = l1.z; y[3] = l1.w;
    a[0] = w.x; a[1] = w.y; a[2] = w.z; a[3] = w.w;
  }
  static __device__ __forceinline__ void st(float* gloc2, float* gawp, const float (&gx)[4],
                                            const float (&gy)[4], const float (&ga)[4]) {
    __stcs(reinterpret_cast<float4*>(gloc2), make_float4(gx[0], gy[0], gx[1], gy[1]));
    __stcs(reinterpret_cast<float4*>(gloc2) + 1, make_float4(gx[2], gy[2], gx[3], gy[3]));
    __stcs(reinterpret_cast<float4*>(gawp), make_float4(ga[0], ga[1], ga[2], ga[3]));
  }
};
template <> struct WinPts<__nv_bfloat16> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* loc2, const __nv_bfloat16* awp,
                                            float (&x)[4], float (&y)[4], float (&a)[4]) {
    const uint4 l = __ldg(reinterpret_cast<const uint4*>(loc2));
    const uint2 w = __ldg(reinterpret_cast<const uint2*>(awp));
    const uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      x[p] = __uint_as_float(lw[p] << 16);
      y[p] = __uint_as_float(lw[p] & 0xffff0000u);
    }
    a[0] = __uint_as_float(w.x << 16); a[1] = __uint_as_float(w.x & 0xffff0000u);
    a[2] = __uint_as_float(w.y << 16); a[3] = __uint_as_float(w.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void ld_stream(const __nv_bfloat16* loc2, const __nv_bfloat16* awp,
                                                   float (&x)[4], float (&y)[4], float (&a)[4]) {
    const uint4 l = __ldcs(reinterpret_cast<const uint4*>(loc2));
    const uint2 w = __ldcs(reinterpret_cast<const uint2*>(awp));
    const uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      x[p] = __uint_as_float(lw[p] << 16);
      y[p] = __uint_as_float(lw[p] & 0xffff0000u);
    }
    a[0] = __uint_as_float(w.x << 16); a[1] = __uint_as_float(w.x & 0xffff0000u);
    a[2] = __uint_as_float(w.y << 16); a[3] = __uint_as_float(w.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void ld_smem(const __nv_bfloat16* loc2, const __nv_bfloat16* awp,
                                                 float (&x)[4], float (&y)[4], float (&a)[4]) {
    const uint4 l = *reinterpret_cast<const uint4*>(loc2);
    const uint2 w = *reinterpret_cast<const uint2*>(awp);
    const uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      x[p] = __uint_as_float(lw[p] << 16);
      y[p] = __uint_as_float(lw[p] & 0xffff0000u);
    }
    a[0] = __uint_as_float(w.x << 16); a[1] = __uint_as_float(w.x & 0xffff0000u);
    a[2] = __uint_as_float(w.y << 16); a[3] = __uint_as_float(w.y & 0xffff0000u);
  }
  static __device__ __forceinline__ uint32_t pack(float lo, float hi) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* gloc2, __nv_bfloat16* gawp,
                                            const float (&gx)[4], const float (&gy)[4],
                                            const float (&ga)[4]) {
    __stcs(reinterpret_cast<uint4*>(gloc2), make_uint4(pack(gx[0], gy[0]), pack(gx[1], gy[1]),
                                                       pack(gx[2], gy[2]), pack(gx[3], gy[3])));
    __stcs(reinterpret_cast<uint2*>(gawp), make_uint2(pack(ga[0], ga[1]), pack(ga[2], ga[3])));
  }
};

// base + 32-bit byte offset in one IMAD.WIDE
__device__ __forceinline__ const char* win_addr(const char* base, unsigned off) {
  unsigned long long r;
  asm("mad.wide.u32 %0, %1, 1, %2;" : "=l"(r) : "r"(off), "l"(base));
  return reinterpret_cast<const char*>(r);
}

// 16-byte asynchronous global -> shared copy (LDGSTS), L2-only caching
__device__ __forceinline__ void win_cp16(void* smem_dst, const void* gsrc) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void win_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void win_cp_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Stage (H, W, start) per level; returns (to every thread) whether all levels
// have H == 1.  One barrier.
__device__ __forceinline__ bool win_stage_levels(const int64_t* __restrict__ shapes,
                                                 const int64_t* __restrict__ lsi, int L,
                                                 int* s_lvl) {
  bool one = true;
  if ((int)threadIdx.x < L) {
    const int l = threadIdx.x;
    const int H = (int)shapes[2 * l];
    one = H == 1;
    s_lvl[3 * l] = H;
    s_lvl[3 * l + 1] = (int)shapes[2 * l + 1];
    s_lvl[3 * l + 2] = (int)lsi[l];
  }
  return __syncthreads_and(one) != 0;
}

// ---- work distribution ------------------------------------------------------
// A UNIT = the 32 / G blocks of one warp (one per G-lane group): 16 consecutive
// queries of one (batch, head) for G = 8, 8 for G = 16.  Warp w of the (persistent) grid takes units w, w + W,
// w + 2W, ... (W = warps in the grid), so at any moment the resident warps work
// on neighbouring units -- the same few (b, m) slabs, which then stay in L2 --
// exactly as a non-persistent launch would sweep them.
struct WinUnit {
  int u, b, m, qc;                       // unit index and its (batch, head, 16-query chunk)
  // first query of the unit; qpu = queries per unit = (32 / G) blocks of kWinQ
  __device__ __forceinline__ int q0(int qpu) const { return qc * qpu; }
};
// The stride W between a warp's units, split once into (chunks, heads, batches):
// stepping a unit is then three adds with carry instead of two divisions.
struct WinStride { int du, dq, dm, db; };

__device__ __forceinline__ WinUnit win_first_unit(int u, int nqc, int M) {
  WinUnit w;
  const int bm = u / nqc;
  w.u = u;
  w.qc = u - bm * nqc;
  w.b = bm / M;
  w.m = bm - w.b * M;
  return w;
}
__device__ __forceinline__ WinStride win_stride(int du, int nqc, int M) {
  WinStride s;
  const int t = du / nqc;
  s.du = du;
  s.dq = du - t * nqc;
  s.db = t / M;
  s.dm = t - s.db * M;
  return s;
}
__device__ __forceinline__ WinUnit win_next_unit(const WinUnit& w, const WinStride& s, int nqc, int M) {
  WinUnit n;
  n.u = w.u + s.du;
  int qc = w.qc + s.dq;
  const int cq = qc >= nqc;
  n.qc = qc - (cq ? nqc : 0);
  int m = w.m + s.dm + cq;
  const int cm = m >= M;
  n.m = m - (cm ? M : 0);
  n.b = w.b + s.db + cm;
  return n;
}

// Shared memory of one group (block of 4 queries):
//   float4   Wt[kWinRows]            weights, later (backward) the dots G
//   unsigned rowoff[kWinRows]        byte offset of the slab row behind packed row t
//   T        pts[NPB][4][LP*3]       staged sampling points of the block (loc then aw per query)
//   T        gout[4][D]              (backward) staged grad_out rows
template <typename T, int NPB, bool BWD, int D>
struct WinSmem {
  static constexpr int kPtsBytes = 4 * (kWinMaxL * kWinP * 3) * (int)sizeof(T);   // per buffer
  static constexpr int kGoutBytes = BWD ? 4 * D * (int)sizeof(T) : 0;
  static constexpr int kPlanBytes = kWinRows * 20;
  static constexpr int kGroupBytes = kPlanBytes + NPB * kPtsBytes + kGoutBytes + 16;  // +16: bank skew
};

// cp.async the sampling points of block (b, q0.., m) into `buf`: loc of the 4
// queries, then their attention weights; queries >= Lq are skipped.  A query's
// loc is at most 8 and its aw at most 4 16-byte pieces (L <= 4, P = 4), so lane j
// copies piece j of each.
template <typename T>
__device__ __forceinline__ void win_fetch_points(const T* __restrict__ loc, const T* __restrict__ aw,
                                                 const Dims& d, int b, int m, int q0, char* buf,
                                                 int j) {
  const int LP = d.L * kWinP;
  const int loc_chunks = LP * 2 * (int)sizeof(T) / 16;
  const int aw_chunks = LP * (int)sizeof(T) / 16;
  const size_t grp = ((size_t)b * d.Lq + q0) * d.M + m;
  const char* lb = reinterpret_cast<const char*>(loc + grp * LP * 2) + j * 16;
  const char* ab = reinterpret_cast<const char*>(aw + grp * LP) + j * 16;
  const unsigned lstride = (unsigned)(d.M * LP * 2) * (unsigned)sizeof(T);   // next query, same head
  char* ldst = buf + j * 16;
  char* adst = buf + (4 * loc_chunks + j) * 16;
#pragma unroll
  for (int q = 0; q < kWinQ; ++q) {
    if (q0 + q < d.Lq) {
      if (j < loc_chunks) win_cp16(ldst + q * loc_chunks * 16, win_addr(lb, q * lstride));
      if (j < aw_chunks) win_cp16(adst + q * aw_chunks * 16, win_addr(ab, q * (lstride >> 1)));
    }
  }
}
template <typename T>
__device__ __forceinline__ const T* win_pts_loc(const char* pts, int L, int qi, int l) {
  return reinterpret_cast<const T*>(pts) + (qi * L * kWinP + l * kWinP) * 2;
}
template <typename T>
__device__ __forceinline__ const T* win_pts_aw(const char* pts, int L, int qi, int l) {
  return reinterpret_cast<const T*>(pts) + 4 * L * kWinP * 2 + qi * L * kWinP + l * kWinP;
}

// Steps 1 + 2 of the header comment, executed by all 32 lanes of the warp, with
// the block's points already staged in shared memory (`pts`).
// Returns ok = false when the block must use the per-tap path (its packed
// windows exceed kWinCap); such a block, and one without queries, takes part
// in the row loop with zero rows.  Outputs: tb[i] = packed position minus slab
// row for the lane's i-th level, rt = this block's packed rows, wmax = the
// largest rt of the warp's blocks (the warp-uniform trip count).  Wt is zero
// and rowoff (BYTE offsets of the slab rows) repeats the last row up to
// round4(wmax) + 4, so the row loops can run past rt.
template <typename T, int G>
__device__ __forceinline__ bool win_build(const char* pts, const Dims& d, bool have_q,
                                          const int* s_lvl, char* s_grp, int j, int gbase,
                                          int (&tb)[(kWinMaxL * 4 + G - 1) / G], int& rt,
                                          int& wmax) {
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int NH = G / 4;                        // level classes
  constexpr int LPL = (kWinMaxL + NH - 1) / NH;    // levels owned per lane
  constexpr int BIG = 0x3fffffff;
  const int qi = j & 3, h = j >> 2;
  float4* Wt = reinterpret_cast<float4*>(s_grp);
  float* Wf = reinterpret_cast<float*>(s_grp);
  unsigned* rowoff = reinterpret_cast<unsigned*>(s_grp + 16 * kWinRows);

  int r0[LPL][4];
  float w0[LPL][4], w1[LPL][4];
  int lo[LPL], hi[LPL];
#pragma unroll
  for (int i = 0; i < LPL; ++i) {
    const int l = h + i * NH;
    const bool active = have_q && l < d.L;
    const int ls = l < d.L ? l : 0;
    float x[4], y[4], a[4];
    WinPts<T>::ld_smem(win_pts_loc<T>(pts, d.L, qi, ls), win_pts_aw<T>(pts, d.L, qi, ls), x, y, a);
    const int W = s_lvl[3 * ls + 1], start = s_lvl[3 * ls + 2];
    lo[i] = BIG; hi[i] = -BIG;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      WinCoef c;
      bool live, v0, v1;
      win_point<false>(x[p], y[p], a[p], W, start, d.border != 0, active, r0[i][p], w0[i][p],
                       w1[i][p], live, v0, v1, c);
      lo[i] = min(lo[i], live ? r0[i][p] : BIG);
      hi[i] = max(hi[i], live ? r0[i][p] + 1 : -BIG);
    }
    // a live point has -1 <= w_low <= W - 1: clip the pair to the map
    lo[i] = max(lo[i], start);
    hi[i] = min(hi[i], start + W - 1);
  }
  // agree per level over the 4 lanes (qi = 0..3) of the class
#pragma unroll
  for (int s = 1; s <= 2; s <<= 1)
#pragma unroll
    for (int i = 0; i < LPL; ++i) {
      lo[i] = min(lo[i], __shfl_xor_sync(FULL, lo[i], s));
      hi[i] = max(hi[i], __shfl_xor_sync(FULL, hi[i], s));
    }
  // every lane learns all levels; pack them back to back
  int plo[kWinMaxL], poff[kWinMaxL];
  int rtot = 0, last_row = 0;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int src = gbase + (l % NH) * 4;
    const int llo = __shfl_sync(FULL, lo[l / NH], src);
    const int lhi = __shfl_sync(FULL, hi[l / NH], src);
    const int rows = max(lhi - llo + 1, 0);
    plo[l] = llo;
    poff[l] = rtot;
    rtot += rows;
    if (rows > 0) last_row = lhi;
  }
  const bool ok = rtot <= kWinCap;
  rt = ok ? rtot : 0;
  if (!ok) last_row = 0;
  wmax = __reduce_max_sync(FULL, rt);

  const unsigned row_bytes = (unsigned)(d.M * d.D) * (unsigned)sizeof(T);
  const int nfill = ((wmax + kWinFillPad - 1) & ~(kWinFillPad - 1)) + kWinFillPad;
  for (int t = j; t < nfill; t += G) {
    Wt[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    int base = plo[0] - poff[0];
#pragma unroll
    for (int l = 1; l < kWinMaxL; ++l)
      if (t >= poff[l]) base = plo[l] - poff[l];
    rowoff[t] = (unsigned)(t < rt ? base + t : last_row) * row_bytes;
  }
#pragma unroll
  for (int i = 0; i < LPL; ++i) {
    const int l = h + i * NH;
    int o = poff[0] - plo[0];
#pragma unroll
    for (int ll = 1; ll < kWinMaxL; ++ll)
      if (l == ll) o = poff[ll] - plo[ll];
    tb[i] = o;
  }
  __syncwarp();
  // scatter: column (level, qi) belongs to this lane alone.  A zero weight needs
  // no store (and a tap outside the map has weight 0), so no validity flags.
  if (ok) {
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      float c0[LPL], c1[LPL];
      int t0[LPL];
#pragma unroll
      for (int i = 0; i < LPL; ++i) {
        t0[i] = (r0[i][p] + tb[i]) * 4 + qi;
        c0[i] = w0[i][p] != 0.f ? Wf[t0[i]] : 0.f;
        c1[i] = w1[i][p] != 0.f ? Wf[t0[i] + 4] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < LPL; ++i) {
        if (w0[i][p] != 0.f) Wf[t0[i]] = c0[i] + w0[i][p];
        if (w1[i][p] != 0.f) Wf[t0[i] + 4] = c1[i] + w1[i][p];
      }
    }
  }
  __syncwarp();
  return ok;
}

// One value row (this lane's NU 16-byte units) as loaded: kept raw until used.
// The load is predicated WITHOUT zero-fill: a skipped load leaves the previous
// (finite) contents, which only ever meet zero weights.
template <typename T, int NU, int G, typename U = Unit<T>>
__device__ __forceinline__ void win_load_row(const char* vbb, unsigned off, bool pred,
                                             typename U::Raw (&buf)[NU][U::RW]) {
  const char* p = win_addr(vbb, off);
#pragma unroll
  for (int u = 0; u < NU; ++u) U::ldg_keep(p + u * (G * U::BYTES), pred, buf[u]);
}

// The first kWinStages rows of a plan: rowoff always holds a valid row there
// (win_build pads with the block's last row, or row 0 of the slab), so these
// loads need no predicate and the buffers no initial value.
template <typename T, int NU, int G, typename U = Unit<T>>
__device__ __forceinline__ void win_load_row_first(const char* vbb, unsigned off,
                                                   typename U::Raw (&buf)[NU][U::RW]) {
  const char* p = win_addr(vbb, off);
#pragma unroll
  for (int u = 0; u < NU; ++u) U::ldg_raw(p + u * (G * U::BYTES), buf[u]);
}

constexpr int kWinStages = 4;   // value rows in flight per lane (8 measured the same: 53.7 us at ENC)

// ------------------------------- forward ------------------------------------
// Persistent CTAs of kWinThreads threads; each warp loops over units.  While a
// block is processed the points of the warp's next unit travel to shared memory
// (cp.async), so only the very first unit of a warp waits on HBM for them.
template <typename T, int NU, int G>
__global__ void __launch_bounds__(kWinThreads)
msda_fwd_win_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                    const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                    const T* __restrict__ aw, T* __restrict__ out, Dims d, int nqc,
                    int units) {
  using Raw = typename Unit<T>::Raw;
  using SM = WinSmem<T, 1, false, 0>;
  constexpr int E = Unit<T>::E;
  constexpr int NS = kWinStages;
  constexpr int LPL = (kWinMaxL * 4 + G - 1) / G;
  __shared__ int s_lvl[3 * kWinMaxL];
  extern __shared__ __align__(16) char s_dyn[];
  const bool one_d = win_stage_levels(shapes, lsi, d.L, s_lvl);
  const int ustep = gridDim.x * (kWinThreads / 32);
  const int uend = units;

  const int gi = threadIdx.x / G;
  const int j = threadIdx.x % G;
  const int lane = threadIdx.x & 31;
  constexpr int QPU = (32 / G) * kWinQ;     // queries per unit
  const int gq = (gi % (32 / G)) * kWinQ;   // this group's query offset inside a unit
  char* s_grp = s_dyn + gi * SM::kGroupBytes;
  char* pts = s_grp + SM::kPlanBytes;
  const float4* Wt = reinterpret_cast<const float4*>(s_grp);
  const unsigned* rowoff = reinterpret_cast<const unsigned*>(s_grp + 16 * kWinRows);

  const WinStride ust = win_stride(ustep, nqc, d.M);
  WinUnit w = win_first_unit(blockIdx.x * (kWinThreads / 32) + (threadIdx.x >> 5), nqc, d.M);
  if (!one_d) {
    for (; w.u < uend; w = win_next_unit(w, ust, nqc, d.M)) {
      const int q0 = w.q0(QPU) + gq;
      fwd_body<T, NU, G, 4>(value, loc, aw, out, d, w.b, w.m, q0, min(q0 + kWinQ, d.Lq), 1, s_lvl,
                            reinterpret_cast<TapMeta<4>*>(s_grp));
      __syncwarp();
    }
    return;
  }
  if (w.u < uend) win_fetch_points<T>(loc, aw, d, w.b, w.m, w.q0(QPU) + gq, pts, j);
  win_cp_commit();
  while (w.u < uend) {
    const int b = w.b, m = w.m, q0 = w.q0(QPU) + gq;
    const int nq = max(min(kWinQ, d.Lq - q0), 0);
    const bool have_q = (j & 3) < nq;
    win_cp_wait_all();
    __syncwarp();
    int tb[LPL], rt, wmax;
    const bool ok = win_build<T, G>(pts, d, have_q, s_lvl, s_grp, j, lane & ~(G - 1), tb, rt, wmax);
    // win_build ends with a warp barrier: the staged points are consumed, the
    // buffer can take the next unit's
    const WinUnit wn = win_next_unit(w, ust, nqc, d.M);
    if (wn.u < uend) win_fetch_points<T>(loc, aw, d, wn.b, wn.m, wn.q0(QPU) + gq, pts, j);
    win_cp_commit();

    const char* vbb =
        reinterpret_cast<const char*>(value + ((size_t)b * d.S * d.M + m) * d.D) + j * 16;
    float2 acc[kWinQ][NU][E / 2];
#pragma unroll
    for (int q = 0; q < kWinQ; ++q)
#pragma unroll
      for (int uu = 0; uu < NU; ++uu)
#pragma unroll
        for (int c = 0; c < E / 2; ++c) acc[q][uu][c] = make_float2(0.f, 0.f);

    // NS rows in flight per lane: row t+NS+s is requested as soon as row t+s is consumed
    const int rl = rt;
    Raw buf[NS][NU][4];
    {
      unsigned ro[NS];
#pragma unroll
      for (int s = 0; s < NS; s += 4) {
        const uint4 r4 = *reinterpret_cast<const uint4*>(rowoff + s);
        ro[s] = r4.x; ro[s + 1] = r4.y; ro[s + 2] = r4.z; ro[s + 3] = r4.w;
      }
#pragma unroll
      for (int s = 0; s < NS; ++s) win_load_row_first<T, NU, G>(vbb, ro[s], buf[s]);
    }
    for (int t = 0; t < wmax; t += NS) {
      unsigned ro[NS];
#pragma unroll
      for (int s = 0; s < NS; s += 4) {
        const uint4 r4 = *reinterpret_cast<const uint4*>(rowoff + t + NS + s);
        ro[s] = r4.x; ro[s + 1] = r4.y; ro[s + 2] = r4.z; ro[s + 3] = r4.w;
      }
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const float4 wv = Wt[t + s];
        const float wq[kWinQ] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int uu = 0; uu < NU; ++uu) {
          float v[E];
          Unit<T>::unpack_raw(buf[s][uu], v);
#pragma unroll
          for (int q = 0; q < kWinQ; ++q) {
            const float2 w2 = make_float2(wq[q], wq[q]);
#pragma unroll
            for (int c = 0; c < E / 2; ++c)
              acc[q][uu][c] = __ffma2_rn(w2, make_float2(v[2 * c], v[2 * c + 1]), acc[q][uu][c]);
          }
        }
        win_load_row<T, NU, G>(vbb, ro[s], t + NS + s < rl, buf[s]);
      }
    }
    if (ok) {
#pragma unroll
      for (int q = 0; q < kWinQ; ++q) {
        if (q < nq) {
          T* og = out + (((size_t)b * d.Lq + q0 + q) * d.M + m) * d.D + j * E;
#pragma unroll
          for (int uu = 0; uu < NU; ++uu) {
            float o[E];
#pragma unroll
            for (int c = 0; c < E / 2; ++c) { o[2 * c] = acc[q][uu][c].x; o[2 * c + 1] = acc[q][uu][c].y; }
            Unit<T>::st_stream(og + uu * (G * E), o);
          }
        }
      }
    } else {
      // the plan area doubles as the per-tap body's scratch
      fwd_body<T, NU, G, 2>(value, loc, aw, out, d, b, m, q0, q0 + nq, 1, s_lvl,
                            reinterpret_cast<TapMeta<2>*>(s_grp));
    }
    __syncwarp();
    w = wn;
  }
}

// ------------------------------- backward -----------------------------------
// One step = the two packed rows t, t+1: grad_value rows leave as 128-bit reds,
// the 2 x 4 dots <grad_out[q], value row> are transpose-reduced over the 8
// lanes; lane (qi, h) ends with the total for (row t+h, query qi).
template <typename T, int NU, int G, typename GV = float, typename U = Unit<T>>
__device__ __forceinline__ void bwd_win_step(const float4* Wt, float* Gs, char* gvbb, unsigned ro0,
                                             unsigned ro1, int t, int rt, int j,
                                             const typename U::Raw (&b0)[NU][U::RW],
                                             const typename U::Raw (&b1)[NU][U::RW],
                                             const float2 (&go)[kWinQ][NU][U::E / 2],
                                             float gv_scale) {
  constexpr int E = U::E;
  constexpr unsigned FULL = 0xffffffffu;
  constexpr unsigned kAccScale = (unsigned)(sizeof(GV) / sizeof(T));  // accumulation image of a T slab
  float x[2 * kWinQ];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const float4 w = Wt[t + r];
    const float wq[kWinQ] = {w.x, w.y, w.z, w.w};
    GV* gp = reinterpret_cast<GV*>(
        const_cast<char*>(win_addr(gvbb, (r == 0 ? ro0 : ro1) * kAccScale)));
    const bool real = t + r < rt;
    float2 s[kWinQ];
#pragma unroll
    for (int q = 0; q < kWinQ; ++q) s[q] = make_float2(0.f, 0.f);
#pragma unroll
    for (int u = 0; u < NU; ++u) {
      // grad_value row: sum_q Wt[t][q] * grad_out[q]
      float2 dv[E / 2];
#pragma unroll
      for (int c = 0; c < E / 2; ++c) dv[c] = make_float2(0.f, 0.f);
#pragma unroll
      for (int q = 0; q < kWinQ; ++q) {
        const float2 w2 = make_float2(wq[q], wq[q]);
#pragma unroll
        for (int c = 0; c < E / 2; ++c) dv[c] = __ffma2_rn(w2, go[q][u][c], dv[c]);
      }
#pragma unroll
      for (int c = 0; c < E / 2; c += 2)
        GvRed<GV>::add4(gp + u * (G * E) + 2 * c, real, dv[c].x, dv[c].y, dv[c + 1].x, dv[c + 1].y,
                        gv_scale);
      // this lane's share of <grad_out[q], value row>
      float v[E];
      U::unpack_raw(r == 0 ? b0[u] : b1[u], v);
#pragma unroll
      for (int q = 0; q < kWinQ; ++q)
#pragma unroll
        for (int c = 0; c < E / 2; ++c)
          s[q] = __ffma2_rn(go[q][u][c], make_float2(v[2 * c], v[2 * c + 1]), s[q]);
    }
#pragma unroll
    for (int q = 0; q < kWinQ; ++q) x[r * kWinQ + q] = s[q].x + s[q].y;
  }
  const float tot = transpose_reduce<8>(x, j, FULL);
  __syncwarp();
  Gs[t * 4 + j] = tot;   // (row t + (j >> 2), query j & 3)
}

template <typename T, int NU, int G, typename GV = float, typename U = Unit<T>>
__global__ void __launch_bounds__(kWinThreads, 4)
msda_bwd_win_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                    const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                    const T* __restrict__ aw, const T* __restrict__ gout,
                    GV* __restrict__ gvalue, T* __restrict__ gloc, T* __restrict__ gaw, Dims d,
                    int nqc, int units, const DetHeader* __restrict__ det) {
  const float gv_scale = det ? (float)det->scale : 1.f;   // fixed-point accumulation (msda_det.cuh)
  static_assert(G == 8, "two rows x four queries are transpose-reduced over 8 lanes");
  using Raw = typename U::Raw;
  constexpr int E = U::E;
  using SM = WinSmem<T, 1, false, 0>;
  constexpr int NH = G / 4;
  constexpr int NS = 4;
  constexpr int LPL = (kWinMaxL * 4 + G - 1) / G;
  __shared__ int s_lvl[3 * kWinMaxL];
  extern __shared__ __align__(16) char s_dyn[];
  const bool one_d = win_stage_levels(shapes, lsi, d.L, s_lvl);
  const int ustep = gridDim.x * (kWinThreads / 32);
  const int uend = units;

  const int gi = threadIdx.x / G;
  const int j = threadIdx.x % G;
  const int lane = threadIdx.x & 31;
  constexpr int QPU = (32 / G) * kWinQ;
  const int gq = (gi % (32 / G)) * kWinQ;
  const int qi = j & 3, h = j >> 2;
  char* s_grp = s_dyn + gi * SM::kGroupBytes;
  char* pts_buf = s_grp + SM::kPlanBytes;   // one staging buffer: less shared memory = more L1
  const float4* Wt = reinterpret_cast<const float4*>(s_grp);
  float* Gs = reinterpret_cast<float*>(s_grp);  // the dots G[t][qi] overwrite Wt[t][qi] step by step
  const unsigned* rowoff = reinterpret_cast<const unsigned*>(s_grp + 16 * kWinRows);

  const WinStride ust = win_stride(ustep, nqc, d.M);
  WinUnit w = win_first_unit(blockIdx.x * (kWinThreads / 32) + (threadIdx.x >> 5), nqc, d.M);
  if (!one_d) {
    for (; w.u < uend; w = win_next_unit(w, ust, nqc, d.M)) {
      const int q0 = w.q0(QPU) + gq;
      bwd_body<T, NU, G, 4, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, w.b, w.m, q0,
                                min(q0 + kWinQ, d.Lq), 1, s_lvl, reinterpret_cast<TapMeta<4>*>(s_grp),
                                gv_scale);
      __syncwarp();
    }
    return;
  }
  if (w.u < uend) win_fetch_points<T>(loc, aw, d, w.b, w.m, w.q0(QPU) + gq, pts_buf, j);
  win_cp_commit();
  while (w.u < uend) {
    const int b = w.b, m = w.m, q0 = w.q0(QPU) + gq;
    const int nq = max(min(kWinQ, d.Lq - q0), 0);
    const bool have_q = qi < nq;
    const char* pts = pts_buf;
    // grad_out of the 4 queries (this lane's channels): requested first, used
    // after the window build, whose work covers their way from HBM
    Raw gor[kWinQ][NU][U::RW];
    {
      const char* gb = reinterpret_cast<const char*>(
          gout + (((size_t)b * d.Lq + q0) * d.M + m) * d.D) + j * U::BYTES;
      const unsigned qstride = (unsigned)(d.M * d.D) * (unsigned)sizeof(T);
#pragma unroll
      for (int q = 0; q < kWinQ; ++q)
#pragma unroll
        for (int uu = 0; uu < NU; ++uu) {
#pragma unroll
          for (int k = 0; k < U::RW; ++k) gor[q][uu][k] = Raw(0);
          U::ldg_keep_stream(win_addr(gb, q * qstride) + uu * (G * U::BYTES), q < nq, gor[q][uu]);
        }
    }
    win_cp_wait_all();
    __syncwarp();
    int tb[LPL], rt, wmax;
    const bool ok = win_build<T, G>(pts, d, have_q, s_lvl, s_grp, j, lane & ~(G - 1), tb, rt, wmax);
    // win_build has consumed the staged points: the buffer takes the next unit's; the epilogue
    // reads this unit's points again from global memory (L2 hits)
    const WinUnit wn = win_next_unit(w, ust, nqc, d.M);
    if (wn.u < uend)
      win_fetch_points<T>(loc, aw, d, wn.b, wn.m, wn.q0(QPU) + gq, pts_buf, j);
    win_cp_commit();
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

    const size_t voff = ((size_t)b * d.S * d.M + m) * d.D;
    const char* vbb = reinterpret_cast<const char*>(value + voff) + j * U::BYTES;
    char* gvbb = reinterpret_cast<char*>(gvalue + voff + j * E);

    // NS rows in flight per lane
    const int rl = rt;
    Raw buf[NS][NU][U::RW];
    const int rr = rt;                      // rows whose grad_value reds are issued
    uint4 rc = *reinterpret_cast<const uint4*>(rowoff);
    win_load_row_first<T, NU, G, U>(vbb, rc.x, buf[0]);
    win_load_row_first<T, NU, G, U>(vbb, rc.y, buf[1]);
    win_load_row_first<T, NU, G, U>(vbb, rc.z, buf[2]);
    win_load_row_first<T, NU, G, U>(vbb, rc.w, buf[3]);
    for (int t = 0; t < wmax; t += NS) {
      const uint4 rn = *reinterpret_cast<const uint4*>(rowoff + t + NS);
      bwd_win_step<T, NU, G, GV, U>(Wt, Gs, gvbb, rc.x, rc.y, t, rr, j, buf[0], buf[1], go, gv_scale);
      win_load_row<T, NU, G, U>(vbb, rn.x, t + NS + 0 < rl, buf[0]);
      win_load_row<T, NU, G, U>(vbb, rn.y, t + NS + 1 < rl, buf[1]);
      bwd_win_step<T, NU, G, GV, U>(Wt, Gs, gvbb, rc.z, rc.w, t + 2, rr, j, buf[2], buf[3], go, gv_scale);
      win_load_row<T, NU, G, U>(vbb, rn.z, t + NS + 2 < rl, buf[2]);
      win_load_row<T, NU, G, U>(vbb, rn.w, t + NS + 3 < rl, buf[3]);
      rc = rn;
    }
    __syncwarp();

    if (ok) {
      // per-point gradients: the staged points again, the geometry again, the dots
      const int LP = d.L * kWinP;
      const size_t grp_me = ((size_t)b * d.Lq + (have_q ? q0 + qi : 0)) * d.M + m;
#pragma unroll
      for (int i = 0; i < LPL; ++i) {
        const int l = h + i * NH;
        const bool active = have_q && l < d.L;
        const int ls = l < d.L ? l : 0;
        float x[4], y[4], a[4];
        // the staging buffer already belongs to the next unit: the points come from L2 again
        WinPts<T>::ld_stream(loc + (grp_me * LP + ls * kWinP) * 2, aw + grp_me * LP + ls * kWinP, x, y, a);
        const int W = s_lvl[3 * ls + 1], start = s_lvl[3 * ls + 2];
        float gx[4], gy[4], ga[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          int r0;
          float w0, w1;
          bool live, v0, v1;
          WinCoef c;
          win_point<true>(x[p], y[p], a[p], W, start, d.border != 0, active, r0, w0, w1, live, v0,
                          v1, c);
          const int t0 = (r0 + tb[i]) * 4 + qi;
          const float g0 = v0 ? Gs[t0] : 0.f;
          const float g1 = v1 ? Gs[t0 + 4] : 0.f;
          ga[p] = fmaf(c.k0, g0, c.k1 * g1);
          gx[p] = c.fx * (g1 - g0);
          gy[p] = fmaf(c.cy0, g0, c.cy1 * g1);
        }
        if (active)
          WinPts<T>::st(gloc + (grp_me * LP + l * kWinP) * 2, gaw + grp_me * LP + l * kWinP, gx, gy, ga);
      }
    } else {
      bwd_body<T, NU, G, 2, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, q0, q0 + nq, 1, s_lvl,
                                reinterpret_cast<TapMeta<2>*>(s_grp), gv_scale);
    }
    __syncwarp();
    w = wn;
  }
}

}  // namespace msda
