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
constexpr int kWinCap = 96;      // packed rows per block, multiple of 4
constexpr int kWinThreads = 128;
// per group: float4 Wt[kWinCap], int rowoff[kWinCap], + 4 floats so that the
// groups of a warp start 4 banks apart
constexpr int kWinGroupFloats = 5 * kWinCap + 4;

struct WinCoef { float k0, k1, fx, cy0, cy1; };

// One sampling point on a (1, W) map: slab row of the low tap, weights (already
// times the attention weight and the vertical factor), which taps are read.
// Same arithmetic as point_geom() with H == 1 (msda_common.cuh).
template <bool BWD>
__device__ __forceinline__ void win_point(float x, float y, float a, int W, int start, bool border,
                                          bool active, int& r0, float& w0, float& w1, bool& v0,
                                          bool& v1, WinCoef& c) {
  float h_im = y - 0.5f;
  float w_im = x * (float)W - 0.5f;
  float sx = (float)W, sy = 1.f;
  bool live;
  if (border) {
    const float wmax = (float)(W - 1);
    if (!(w_im > 0.f)) { w_im = 0.f; sx = 0.f; }
    else if (w_im >= wmax) { w_im = wmax; sx = 0.f; }
    h_im = 0.f; sy = 0.f;
    live = active;
  } else {
    live = active && (h_im > -1.f) && (w_im > -1.f) && (h_im < 1.f) && (w_im < (float)W);
  }
  const float wf = floorf(w_im);
  const int w_low = (int)wf;
  const float lw = w_im - wf, hw = 1.f - lw;
  const bool hl = h_im >= 0.f;                       // row 0 is the LOW row of the vertical pair
  const float fy = hl ? 1.f - h_im : h_im + 1.f;     // hh or lh
  v0 = live && w_low >= 0;
  v1 = live && w_low + 1 <= W - 1;
  r0 = start + w_low;
  const float k0 = fy * hw, k1 = fy * lw;
  w0 = v0 ? k0 * a : 0.f;
  w1 = v1 ? k1 * a : 0.f;
  if (BWD) {
    c.k0 = v0 ? k0 : 0.f;
    c.k1 = v1 ? k1 : 0.f;
    c.fx = live ? sx * a * fy : 0.f;
    const float sg = hl ? -sy : sy;
    c.cy0 = live ? sg * a * hw : 0.f;
    c.cy1 = live ? sg * a * lw : 0.f;
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

// What every lane of a block knows after the window agreement.
struct WinPlan {
  int lo[kWinMaxL];     // first slab row of level l's window
  int off[kWinMaxL];    // packed position of that row
  int rtot;             // packed rows over all levels
};

// Steps 1 + 2 of the header comment.  Returns false when the block must use the
// per-tap path (packed windows exceed kWinCap).  On success Wt / rowoff are
// built and visible to the whole group.
template <typename T, int G>
__device__ __forceinline__ bool win_build(const T* __restrict__ loc, const T* __restrict__ aw,
                                          const Dims& d, size_t grp_me, bool have_q,
                                          const int* s_lvl, float* s_grp, int j, int gbase,
                                          unsigned mask, WinPlan& pl) {
  constexpr int NH = G / 4;                        // level classes
  constexpr int LPL = (kWinMaxL + NH - 1) / NH;    // levels owned per lane
  constexpr int BIG = 0x3fffffff;
  const int qi = j & 3, h = j >> 2;
  const int MD = d.M * d.D;
  const int LP = d.L * kWinP;
  float4* Wt = reinterpret_cast<float4*>(s_grp);
  float* Wf = s_grp;
  int* rowoff = reinterpret_cast<int*>(s_grp + 4 * kWinCap);

  int r0[LPL][4];
  float w0[LPL][4], w1[LPL][4];
  bool v0[LPL][4], v1[LPL][4];
  int lo[LPL], hi[LPL];
#pragma unroll
  for (int i = 0; i < LPL; ++i) {
    const int l = h + i * NH;
    const bool active = have_q && l < d.L;
    const int ls = active ? l : 0;
    float x[4], y[4], a[4];
    WinPts<T>::ld(loc + (grp_me * LP + ls * kWinP) * 2, aw + grp_me * LP + ls * kWinP, x, y, a);
    const int W = s_lvl[3 * ls + 1], start = s_lvl[3 * ls + 2];
    lo[i] = BIG; hi[i] = -BIG;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      WinCoef c;
      win_point<false>(x[p], y[p], a[p], W, start, d.border != 0, active, r0[i][p], w0[i][p],
                       w1[i][p], v0[i][p], v1[i][p], c);
      if (v0[i][p]) { lo[i] = min(lo[i], r0[i][p]); hi[i] = max(hi[i], r0[i][p]); }
      if (v1[i][p]) { lo[i] = min(lo[i], r0[i][p] + 1); hi[i] = max(hi[i], r0[i][p] + 1); }
    }
  }
  // agree per level over the 4 lanes (qi = 0..3) of the class
#pragma unroll
  for (int s = 1; s <= 2; s <<= 1)
#pragma unroll
    for (int i = 0; i < LPL; ++i) {
      lo[i] = min(lo[i], __shfl_xor_sync(mask, lo[i], s));
      hi[i] = max(hi[i], __shfl_xor_sync(mask, hi[i], s));
    }
  // every lane learns all levels; pack them back to back
  int rtot = 0;
#pragma unroll
  for (int l = 0; l < kWinMaxL; ++l) {
    const int src = gbase + (l % NH) * 4;
    const int llo = __shfl_sync(mask, lo[l / NH], src);
    const int lhi = __shfl_sync(mask, hi[l / NH], src);
    const int rows = max(lhi - llo + 1, 0);
    pl.lo[l] = llo;
    pl.off[l] = rtot;
    rtot += rows;
  }
  pl.rtot = rtot;
  if (rtot > kWinCap) return false;

  // zero Wt, fill rowoff (padding rows up to a multiple of 4: weight 0, offset -1)
  const int rpad = (rtot + 3) & ~3;
  for (int t = j; t < rpad; t += G) {
    Wt[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    int base = pl.lo[0] - pl.off[0];
#pragma unroll
    for (int l = 1; l < kWinMaxL; ++l)
      if (t >= pl.off[l]) base = pl.lo[l] - pl.off[l];
    rowoff[t] = t < rtot ? (base + t) * MD : -1;
  }
  __syncwarp(mask);
  // scatter: column (level, qi) belongs to this lane alone
  int tb[LPL];
#pragma unroll
  for (int i = 0; i < LPL; ++i) {
    const int l = h + i * NH;
    int o = pl.off[0] - pl.lo[0];
#pragma unroll
    for (int ll = 1; ll < kWinMaxL; ++ll)
      if (l == ll) o = pl.off[ll] - pl.lo[ll];
    tb[i] = o;
  }
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    float c0[LPL], c1[LPL];
    int t0[LPL];
#pragma unroll
    for (int i = 0; i < LPL; ++i) {
      t0[i] = (v0[i][p] || v1[i][p]) ? (r0[i][p] + tb[i]) * 4 + qi : qi;
      c0[i] = v0[i][p] ? Wf[t0[i]] : 0.f;
      c1[i] = v1[i][p] ? Wf[t0[i] + 4] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < LPL; ++i) {
      if (v0[i][p]) Wf[t0[i]] = c0[i] + w0[i][p];
      if (v1[i][p]) Wf[t0[i] + 4] = c1[i] + w1[i][p];
    }
  }
  __syncwarp(mask);
  return true;
}

// ------------------------------- forward ------------------------------------
template <typename T, int NU, int G>
__device__ __forceinline__ void fwd_win_rows(const T* __restrict__ vb, T* __restrict__ out,
                                             const Dims& d, int b, int m, int q0, int nq,
                                             const float* s_grp, int j, int rtot) {
  constexpr int E = Unit<T>::E;
  constexpr int RB = 4;  // rows per step
  const float4* Wt = reinterpret_cast<const float4*>(s_grp);
  const int* rowoff = reinterpret_cast<const int*>(s_grp + 4 * kWinCap);
  float2 acc[kWinQ][NU][E / 2];
#pragma unroll
  for (int q = 0; q < kWinQ; ++q)
#pragma unroll
    for (int u = 0; u < NU; ++u)
#pragma unroll
      for (int c = 0; c < E / 2; ++c) acc[q][u][c] = make_float2(0.f, 0.f);

  for (int t = 0; t < rtot; t += RB) {
    const int4 ro4 = *reinterpret_cast<const int4*>(rowoff + t);
    const int ro[RB] = {ro4.x, ro4.y, ro4.z, ro4.w};
    float v[RB][NU][E];
#pragma unroll
    for (int r = 0; r < RB; ++r)
#pragma unroll
      for (int u = 0; u < NU; ++u) Unit<T>::ldg(vb + ro[r] + u * (G * E), ro[r] >= 0, v[r][u]);
    float4 w[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) w[r] = Wt[t + r];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const float wq[kWinQ] = {w[r].x, w[r].y, w[r].z, w[r].w};
#pragma unroll
      for (int q = 0; q < kWinQ; ++q) {
        const float2 w2 = make_float2(wq[q], wq[q]);
#pragma unroll
        for (int u = 0; u < NU; ++u)
#pragma unroll
          for (int c = 0; c < E / 2; ++c)
            acc[q][u][c] = __ffma2_rn(w2, make_float2(v[r][u][2 * c], v[r][u][2 * c + 1]), acc[q][u][c]);
      }
    }
  }
#pragma unroll
  for (int q = 0; q < kWinQ; ++q) {
    if (q < nq) {
      T* og = out + (((size_t)b * d.Lq + q0 + q) * d.M + m) * d.D + j * E;
#pragma unroll
      for (int u = 0; u < NU; ++u) {
        float o[E];
#pragma unroll
        for (int c = 0; c < E / 2; ++c) { o[2 * c] = acc[q][u][c].x; o[2 * c + 1] = acc[q][u][c].y; }
        Unit<T>::st_stream(og + u * (G * E), o);
      }
    }
  }
}

// CTA = kWinThreads threads = kWinThreads/G blocks of 4 consecutive queries of one (b, m).
template <typename T, int NU, int G>
__global__ void __launch_bounds__(kWinThreads)
msda_fwd_win_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                    const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                    const T* __restrict__ aw, T* __restrict__ out, Dims d, int nqc) {
  constexpr int E = Unit<T>::E;
  constexpr int NG = kWinThreads / G;
  __shared__ int s_lvl[3 * kMaxLevelsFast];
  __shared__ int s_one_d;
  __shared__ __align__(16) float s_win[NG * kWinGroupFloats];
  stage_levels(shapes, lsi, d.L, s_lvl, &s_one_d);

  const int tile = blockIdx.x;
  const int m = tile % d.M;
  const int qc = (tile / d.M) % nqc;
  const int b = tile / (d.M * nqc);
  const int gi = threadIdx.x / G;
  const int j = threadIdx.x % G;
  const int lane = threadIdx.x & 31;
  const int q0 = (qc * NG + gi) * kWinQ;
  const int nq = min(kWinQ, d.Lq - q0);
  if (nq <= 0) return;
  float* s_grp = s_win + gi * kWinGroupFloats;
  if (!s_one_d) {
    fwd_body<T, NU, G, 4>(value, loc, aw, out, d, b, m, q0, q0 + nq, 1, s_lvl,
                          reinterpret_cast<TapMeta<4>*>(s_grp));
    return;
  }
  const unsigned mask = group_mask<G>(lane);
  const bool have_q = (j & 3) < nq;
  const size_t grp_me = ((size_t)b * d.Lq + q0 + (have_q ? (j & 3) : 0)) * d.M + m;
  WinPlan pl;
  if (!win_build<T, G>(loc, aw, d, grp_me, have_q, s_lvl, s_grp, j, lane & ~(G - 1), mask, pl)) {
    fwd_body<T, NU, G, 2>(value, loc, aw, out, d, b, m, q0, q0 + nq, 1, s_lvl,
                          reinterpret_cast<TapMeta<2>*>(s_grp));
    return;
  }
  const T* vb = value + ((size_t)b * d.S * d.M + m) * d.D + j * E;
  fwd_win_rows<T, NU, G>(vb, out, d, b, m, q0, nq, s_grp, j, pl.rtot);
}

// ------------------------------- backward -----------------------------------
template <typename T, int NU, int G>
__device__ __forceinline__ void bwd_win_block(const T* __restrict__ value,
                                              const T* __restrict__ loc, const T* __restrict__ aw,
                                              const T* __restrict__ gout,
                                              float* __restrict__ gvalue, T* __restrict__ gloc,
                                              T* __restrict__ gaw, const Dims& d, int b, int m,
                                              int q0, int nq, size_t grp_me, bool have_q,
                                              const int* s_lvl, float* s_grp, int j, unsigned mask,
                                              const WinPlan& pl) {
  static_assert(G == 8, "transpose-reduce of 2 rows x 4 queries over 8 lanes");
  constexpr int E = Unit<T>::E;
  constexpr int RB = 2;
  constexpr int NH = G / 4;
  constexpr int LPL = (kWinMaxL + NH - 1) / NH;
  float4* Wt = reinterpret_cast<float4*>(s_grp);
  float* Gs = s_grp;  // the dots G[t][qi] overwrite Wt[t][qi] row by row
  const int* rowoff = reinterpret_cast<const int*>(s_grp + 4 * kWinCap);
  const int qi = j & 3, h = j >> 2;
  const size_t voff = ((size_t)b * d.S * d.M + m) * d.D + j * E;
  const T* vb = value + voff;
  float* gvb = gvalue + voff;

  float2 go[kWinQ][NU][E / 2];
#pragma unroll
  for (int q = 0; q < kWinQ; ++q) {
    const T* gg = gout + (((size_t)b * d.Lq + q0 + min(q, nq - 1)) * d.M + m) * d.D + j * E;
#pragma unroll
    for (int u = 0; u < NU; ++u) {
      float t[E];
      Unit<T>::ld_stream(gg + u * (G * E), t);
#pragma unroll
      for (int c = 0; c < E / 2; ++c)
        go[q][u][c] = q < nq ? make_float2(t[2 * c], t[2 * c + 1]) : make_float2(0.f, 0.f);
    }
  }

  const int rtot = pl.rtot;
  for (int t = 0; t < rtot; t += RB) {
    const int2 ro2 = *reinterpret_cast<const int2*>(rowoff + t);
    const int ro[RB] = {ro2.x, ro2.y};
    float v[RB][NU][E];
#pragma unroll
    for (int r = 0; r < RB; ++r)
#pragma unroll
      for (int u = 0; u < NU; ++u) Unit<T>::ldg(vb + ro[r] + u * (G * E), ro[r] >= 0, v[r][u]);
    float4 w[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) w[r] = Wt[t + r];
    float dots[RB * kWinQ];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const float wq[kWinQ] = {w[r].x, w[r].y, w[r].z, w[r].w};
      // grad_value row: sum_q Wt[t][q] * grad_out[q]
#pragma unroll
      for (int u = 0; u < NU; ++u) {
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
          red_add_f32x4_pred(gvb + ro[r] + u * (G * E) + 2 * c, ro[r] >= 0, dv[c].x, dv[c].y,
                             dv[c + 1].x, dv[c + 1].y);
      }
      // this lane's share of <grad_out[q], value row>
#pragma unroll
      for (int q = 0; q < kWinQ; ++q) {
        float2 s = make_float2(0.f, 0.f);
#pragma unroll
        for (int u = 0; u < NU; ++u)
#pragma unroll
          for (int c = 0; c < E / 2; ++c)
            s = __ffma2_rn(go[q][u][c], make_float2(v[r][u][2 * c], v[r][u][2 * c + 1]), s);
        dots[r * kWinQ + q] = s.x + s.y;
      }
    }
    const float tot = transpose_reduce<G>(dots, j, mask);  // lane j: row t + (j >> 2), query j & 3
    __syncwarp(mask);
    Gs[(t + h) * 4 + qi] = tot;
  }
  __syncwarp(mask);

  // per-point gradients: reload the points, redo the geometry, pick the dots
  const int LP = d.L * kWinP;
#pragma unroll
  for (int i = 0; i < LPL; ++i) {
    const int l = h + i * NH;
    const bool active = have_q && l < d.L;
    const int ls = active ? l : 0;
    float x[4], y[4], a[4];
    WinPts<T>::ld(loc + (grp_me * LP + ls * kWinP) * 2, aw + grp_me * LP + ls * kWinP, x, y, a);
    const int W = s_lvl[3 * ls + 1], start = s_lvl[3 * ls + 2];
    int tb = pl.off[0] - pl.lo[0];
#pragma unroll
    for (int ll = 1; ll < kWinMaxL; ++ll)
      if (l == ll) tb = pl.off[ll] - pl.lo[ll];
    float gx[4], gy[4], ga[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      int r0;
      float w0, w1;
      bool v0, v1;
      WinCoef c;
      win_point<true>(x[p], y[p], a[p], W, start, d.border != 0, active, r0, w0, w1, v0, v1, c);
      const int t0 = (v0 || v1) ? (r0 + tb) * 4 + qi : qi;
      const float g0 = v0 ? Gs[t0] : 0.f;
      const float g1 = v1 ? Gs[t0 + 4] : 0.f;
      ga[p] = fmaf(c.k0, g0, c.k1 * g1);
      gx[p] = c.fx * (g1 - g0);
      gy[p] = fmaf(c.cy0, g0, c.cy1 * g1);
    }
    if (active)
      WinPts<T>::st(gloc + (grp_me * LP + l * kWinP) * 2, gaw + grp_me * LP + l * kWinP, gx, gy, ga);
  }
}

template <typename T, int NU, int G>
__global__ void __launch_bounds__(kWinThreads)
msda_bwd_win_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                    const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                    const T* __restrict__ aw, const T* __restrict__ gout,
                    float* __restrict__ gvalue, T* __restrict__ gloc, T* __restrict__ gaw, Dims d,
                    int nqc) {
  constexpr int NG = kWinThreads / G;
  __shared__ int s_lvl[3 * kMaxLevelsFast];
  __shared__ int s_one_d;
  __shared__ __align__(16) float s_win[NG * kWinGroupFloats];
  stage_levels(shapes, lsi, d.L, s_lvl, &s_one_d);

  const int tile = blockIdx.x;
  const int m = tile % d.M;
  const int qc = (tile / d.M) % nqc;
  const int b = tile / (d.M * nqc);
  const int gi = threadIdx.x / G;
  const int j = threadIdx.x % G;
  const int lane = threadIdx.x & 31;
  const int q0 = (qc * NG + gi) * kWinQ;
  const int nq = min(kWinQ, d.Lq - q0);
  if (nq <= 0) return;
  float* s_grp = s_win + gi * kWinGroupFloats;
  if (!s_one_d) {
    bwd_body<T, NU, G, 4>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, q0, q0 + nq, 1, s_lvl,
                          reinterpret_cast<TapMeta<4>*>(s_grp));
    return;
  }
  const unsigned mask = group_mask<G>(lane);
  const bool have_q = (j & 3) < nq;
  const size_t grp_me = ((size_t)b * d.Lq + q0 + (have_q ? (j & 3) : 0)) * d.M + m;
  WinPlan pl;
  if (!win_build<T, G>(loc, aw, d, grp_me, have_q, s_lvl, s_grp, j, lane & ~(G - 1), mask, pl)) {
    bwd_body<T, NU, G, 2>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, q0, q0 + nq, 1, s_lvl,
                          reinterpret_cast<TapMeta<2>*>(s_grp));
    return;
  }
  bwd_win_block<T, NU, G>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, q0, nq, grp_me, have_q,
                          s_lvl, s_grp, j, mask, pl);
}

}  // namespace msda
