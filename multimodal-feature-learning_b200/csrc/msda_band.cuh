// msda_band.cuh -- "band window" MSDeformAttn kernels for the temporal (H == 1)
// use of the op with query locality (encoder self-attention: consecutive
// queries sample neighbouring rows, models/deformable/
// unimodal_deformable_transformer.py:259-276 builds reference points that are
// monotone in the token index).
//
// A TILE = Q consecutive queries of one (batch, head), owned by LG lanes; lane
// j owns E = 16/sizeof(T) channels of the head row.  Per level:
//   A. each lane computes the geometry of the P points of "its" query (lane i
//      <-> query i) and the tile agrees on the window [rmin, rmax] of rows
//      touched (two shuffle reductions);
//   B. the interpolation * attention weights are scattered into a small
//      shared-memory matrix Wt[row - rmin][query] (column i is private to lane
//      i: no atomics);
//   C. forward : out[q] += Wt[r][q] * value[r] for every row of the window --
//                each value row is loaded ONCE per tile (not once per tap) and
//                the accumulators out[Q][E] live in registers;
//      backward: grad_value[r] = sum_q Wt[r][q] * grad_out[q] is formed in
//                registers and leaves as ONE 128-bit red per row and lane per
//                tile instead of one per tap (2*P*Q of them).
// With Q = 16, P = 4 and the offsets of a freshly initialised module the
// window is ~Q*T_l/T_q + 6 rows, i.e. 3.4 row loads (or reds) per query and
// tile over the four levels instead of 32.
// A level whose window exceeds kBandRows (random locations, a tile straddling
// two query levels, much finer key level) falls back, for that tile and level
// only, to per-tap gathers / reds with identical arithmetic.
// The backward's dot products <grad_out, v_tap> (grad_sampling_loc,
// grad_attn_weight) stay per tap: transpose-reduce over the tile hands query
// i's totals to lane i.
#pragma once
#include "msda_fast.cuh"

namespace msda {

constexpr int kBandRows = 48;  // window capacity (rows) per tile and level

template <int LG> __device__ __forceinline__ int tile_min(int v, unsigned mask) {
#pragma unroll
  for (int s = LG / 2; s >= 1; s >>= 1) v = min(v, __shfl_xor_sync(mask, v, s));
  return v;
}
template <int LG> __device__ __forceinline__ int tile_max(int v, unsigned mask) {
#pragma unroll
  for (int s = LG / 2; s >= 1; s >>= 1) v = max(v, __shfl_xor_sync(mask, v, s));
  return v;
}

// One sampling point as the band kernels keep it: the row of the LOW tap
// (start + w_low; the high tap is r0 + 1), weights already multiplied by the
// attention weight and the vertical factor, 0 for a tap that is not read.
struct BandPoint {
  int r0;
  float w0, w1;
  bool v0, v1;  // tap is read
};
struct BandCoef {  // backward: d(out)/d(aw), d(out)/d(loc) factors per tap
  float cw0, cw1, cx0, cx1, cy0, cy1;
};

// Raw (x, y, attention weight) of the PP points of one (query, level), loaded
// early so that the latency overlaps the previous level's row loop.
template <int PP> struct RawPoints { float x[PP], y[PP], a[PP]; };

template <typename T, int PP>
__device__ __forceinline__ void load_raw(const T* __restrict__ loc2, const T* __restrict__ awp,
                                         bool live, RawPoints<PP>& rp) {
#pragma unroll
  for (int p = 0; p < PP; ++p) { rp.x[p] = rp.y[p] = 0.f; rp.a[p] = 0.f; }
  if (!live) return;
#pragma unroll
  for (int p = 0; p < PP; ++p) LocLoad<T>::ld(loc2 + 2 * p, awp + p, rp.x[p], rp.y[p], rp.a[p]);
}
template <>
__device__ __forceinline__ void load_raw<float, 4>(const float* __restrict__ loc2,
                                                   const float* __restrict__ awp, bool live,
                                                   RawPoints<4>& rp) {
#pragma unroll
  for (int p = 0; p < 4; ++p) { rp.x[p] = rp.y[p] = 0.f; rp.a[p] = 0.f; }
  if (!live) return;
  const float4 l0 = __ldcs(reinterpret_cast<const float4*>(loc2));
  const float4 l1 = __ldcs(reinterpret_cast<const float4*>(loc2) + 1);
  const float4 a = __ldcs(reinterpret_cast<const float4*>(awp));
  rp.x[0] = l0.x; rp.y[0] = l0.y; rp.x[1] = l0.z; rp.y[1] = l0.w;
  rp.x[2] = l1.x; rp.y[2] = l1.y; rp.x[3] = l1.z; rp.y[3] = l1.w;
  rp.a[0] = a.x; rp.a[1] = a.y; rp.a[2] = a.z; rp.a[3] = a.w;
}

template <bool BWD>
__device__ __forceinline__ void band_point(float lx, float ly, float a, bool live_query, int W,
                                           int start, bool border, BandPoint& bp, BandCoef& bc) {
  bp.r0 = 0; bp.w0 = bp.w1 = 0.f; bp.v0 = bp.v1 = false;
  bc.cw0 = bc.cw1 = bc.cx0 = bc.cx1 = bc.cy0 = bc.cy1 = 0.f;
  if (!live_query) return;
  PointGeom<float> g;
  point_geom<float>(lx, ly, 1, W, start, border, g);
  const bool alt = !(g.off[0] >= 0 || g.off[1] >= 0);  // row 0 is the HIGH row of the pair
  const int o0 = alt ? g.off[2] : g.off[0], o1 = alt ? g.off[3] : g.off[1];
  const float k0 = alt ? g.w[2] : g.w[0], k1 = alt ? g.w[3] : g.w[1];
  bp.v0 = o0 >= 0; bp.v1 = o1 >= 0;
  bp.r0 = bp.v0 ? o0 : o1 - 1;
  bp.w0 = bp.v0 ? k0 * a : 0.f;
  bp.w1 = bp.v1 ? k1 * a : 0.f;
  if (BWD) {
    const float hh = 1.f - g.lh, hw = 1.f - g.lw;
    bc.cw0 = k0; bc.cw1 = k1;
    bc.cx0 = (alt ? -g.lh : -hh) * g.sx * a;
    bc.cx1 = (alt ? g.lh : hh) * g.sx * a;
    bc.cy0 = (alt ? hw : -hw) * g.sy * a;
    bc.cy1 = (alt ? g.lw : -g.lw) * g.sy * a;
  }
}

// Phase A+B for one level.  Returns the window [rmin, rmin + rows); rows == 0
// means "nothing to do", rows < 0 means "window too wide: use the per-tap path".
template <int LG, int Q, int PP>
__device__ __forceinline__ void band_build(const BandPoint (&pts)[PP], int li, unsigned mask,
                                           float2* __restrict__ Wt, int& rmin, int& rows) {
  int lo = 0x7fffffff, hi = -0x7fffffff;
#pragma unroll
  for (int p = 0; p < PP; ++p) {
    if (pts[p].v0) { lo = min(lo, pts[p].r0); hi = max(hi, pts[p].r0); }
    if (pts[p].v1) { lo = min(lo, pts[p].r0 + 1); hi = max(hi, pts[p].r0 + 1); }
  }
  lo = tile_min<LG>(lo, mask);
  hi = tile_max<LG>(hi, mask);
  rmin = lo;
  if (hi < lo) { rows = 0; return; }
  rows = hi - lo + 1;
  if (rows > kBandRows) { rows = -1; return; }
  if (li < Q) {  // column li of Wt[rows][Q] is private to this lane: zero it, then scatter
    for (int t = 0; t < rows; ++t) Wt[t * Q + li] = make_float2(0.f, 0.f);
#pragma unroll
    for (int p = 0; p < PP; ++p) {
      const int t = pts[p].r0 - lo;
      if (pts[p].v0) {
        float2 w = Wt[t * Q + li];
        w.x += pts[p].w0; w.y += pts[p].w0;
        Wt[t * Q + li] = w;
      }
      if (pts[p].v1) {
        float2 w = Wt[(t + 1) * Q + li];
        w.x += pts[p].w1; w.y += pts[p].w1;
        Wt[(t + 1) * Q + li] = w;
      }
    }
  }
  __syncwarp(mask);
}

// ------------------------------- forward ------------------------------------
template <typename T, int LG, int Q, int PP>
__device__ __forceinline__ void fwd_band_tile(const T* __restrict__ value,
                                              const T* __restrict__ loc, const T* __restrict__ aw,
                                              T* __restrict__ out, const Dims& d, int b, int m,
                                              int q0, const int* s_lvl, float2* __restrict__ Wt) {
  static_assert(Q == LG, "one query per lane");
  constexpr int E = Unit<T>::E;
  const int lane = threadIdx.x & 31;
  const int li = threadIdx.x % LG;
  const unsigned mask = group_mask<LG>(lane);
  const int lbase = lane & ~(LG - 1);
  const int MD = d.M * d.D;
  const int LP = d.L * PP;
  const T* vb = value + ((size_t)b * d.S * d.M + m) * d.D + li * E;
  const int nq = min(Q, d.Lq - q0);
  const bool live_q = li < nq;
  const size_t grp_me = ((size_t)b * d.Lq + (q0 + (live_q ? li : 0))) * d.M + m;

  float2 acc[Q][E / 2];
#pragma unroll
  for (int q = 0; q < Q; ++q)
#pragma unroll
    for (int c = 0; c < E / 2; ++c) acc[q][c] = make_float2(0.f, 0.f);

  constexpr int RB = 8;  // value rows in flight per batch
  RawPoints<PP> raw_next;
  load_raw<T, PP>(loc + grp_me * LP * 2, aw + grp_me * LP, live_q, raw_next);
  for (int l = 0; l < d.L; ++l) {
    const int W = s_lvl[3 * l + 1], start = s_lvl[3 * l + 2];
    const RawPoints<PP> raw = raw_next;
    if (l + 1 < d.L)
      load_raw<T, PP>(loc + (grp_me * LP + (l + 1) * PP) * 2, aw + grp_me * LP + (l + 1) * PP, live_q,
                      raw_next);
    BandPoint pts[PP];
#pragma unroll
    for (int p = 0; p < PP; ++p) {
      BandCoef bc;
      band_point<false>(raw.x[p], raw.y[p], raw.a[p], live_q, W, start, d.border != 0, pts[p], bc);
    }
    int rmin, rows;
    band_build<LG, Q, PP>(pts, li, mask, Wt, rmin, rows);
    if (rows > 0) {
      const T* vrow = vb + (size_t)rmin * MD;
      for (int rb = 0; rb < rows; rb += RB) {
        float v[RB][E];
#pragma unroll
        for (int k = 0; k < RB; ++k) Unit<T>::ldg(vrow + (size_t)(rb + k) * MD, rb + k < rows, v[k]);
#pragma unroll
        for (int k = 0; k < RB; ++k) {
          if (rb + k < rows) {
            const float4* wp = reinterpret_cast<const float4*>(Wt + (rb + k) * Q);
#pragma unroll
            for (int q = 0; q < Q; q += 2) {
              const float4 w = wp[q / 2];
#pragma unroll
              for (int c = 0; c < E / 2; ++c) {
                const float2 vv = make_float2(v[k][2 * c], v[k][2 * c + 1]);
                acc[q][c] = __ffma2_rn(make_float2(w.x, w.y), vv, acc[q][c]);
                acc[q + 1][c] = __ffma2_rn(make_float2(w.z, w.w), vv, acc[q + 1][c]);
              }
            }
          }
        }
      }
    } else if (rows < 0) {
      // per-tap path: query i's points are broadcast from lane i; two queries per
      // iteration keep 4*PP loads in flight
#pragma unroll 1
      for (int q = 0; q < nq; q += 2) {
        float2 t[2][E / 2];
#pragma unroll
        for (int u = 0; u < 2; ++u)
#pragma unroll
          for (int c = 0; c < E / 2; ++c) t[u][c] = make_float2(0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int src = lbase + min(q + u, Q - 1);
#pragma unroll
          for (int p = 0; p < PP; ++p) {
            const int r0 = __shfl_sync(mask, pts[p].r0, src);
            const float w0 = __shfl_sync(mask, pts[p].w0, src);
            const float w1 = __shfl_sync(mask, pts[p].w1, src);
            const int fl = __shfl_sync(mask, (int)pts[p].v0 | ((int)pts[p].v1 << 1), src);
            float v0[E], v1[E];
            Unit<T>::ldg(vb + (size_t)r0 * MD, (fl & 1) != 0, v0);
            Unit<T>::ldg(vb + (size_t)(r0 + 1) * MD, (fl & 2) != 0, v1);
#pragma unroll
            for (int c = 0; c < E / 2; ++c) {
              t[u][c] = __ffma2_rn(make_float2(w0, w0), make_float2(v0[2 * c], v0[2 * c + 1]), t[u][c]);
              t[u][c] = __ffma2_rn(make_float2(w1, w1), make_float2(v1[2 * c], v1[2 * c + 1]), t[u][c]);
            }
          }
        }
#pragma unroll
        for (int qq = 0; qq < Q; qq += 2)
          if (qq == q) {
#pragma unroll
            for (int c = 0; c < E / 2; ++c) {
              acc[qq][c] = __fadd2_rn(acc[qq][c], t[0][c]);
              if (qq + 1 < nq) acc[qq + 1][c] = __fadd2_rn(acc[qq + 1][c], t[1][c]);
            }
          }
      }
    }
    __syncwarp(mask);
  }
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    if (q < nq) {
      float o[E];
#pragma unroll
      for (int c = 0; c < E / 2; ++c) { o[2 * c] = acc[q][c].x; o[2 * c + 1] = acc[q][c].y; }
      Unit<T>::st_stream(out + (((size_t)b * d.Lq + q0 + q) * d.M + m) * d.D + li * E, o);
    }
  }
}

// CTA = 256 threads = 256/LG tiles of Q consecutive queries of one (b, m).
// Falls back to the 4-tap sparse body when some level has H != 1.
template <typename T, int LG, int Q, int PP, int NU, int G>
__global__ void __launch_bounds__(kFastThreads, 2)
msda_fwd_band_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                     const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                     const T* __restrict__ aw, T* __restrict__ out, Dims d, int nqc) {
  constexpr int kTiles = kFastThreads / LG;
  __shared__ int s_lvl[3 * kMaxLevelsFast];
  __shared__ int s_one_d;
  extern __shared__ __align__(16) unsigned char s_dyn[];  // kTiles * kBandRows * Q float2
  float2* s_W = reinterpret_cast<float2*>(s_dyn);
  stage_levels(shapes, lsi, d.L, s_lvl, &s_one_d);
  const int tile = blockIdx.x;
  const int m = tile % d.M;
  const int qc = (tile / d.M) % nqc;
  const int b = tile / (d.M * nqc);
  const int q_cta = qc * (kTiles * Q);
  if (!s_one_d) {
    fwd_body<T, NU, G, 4>(value, loc, aw, out, d, b, m, q_cta, min(q_cta + kTiles * Q, d.Lq), s_lvl,
                          reinterpret_cast<TapMeta<4>*>(s_W));
    return;
  }
  const int ti = threadIdx.x / LG;
  const int q0 = q_cta + ti * Q;
  if (q0 >= d.Lq) return;
  fwd_band_tile<T, LG, Q, PP>(value, loc, aw, out, d, b, m, q0, s_lvl, s_W + ti * (kBandRows * Q));
}

template <int LG, int Q> constexpr size_t band_smem_bytes() {
  return (size_t)(kFastThreads / LG) * kBandRows * Q * sizeof(float2);
}

}  // namespace msda
