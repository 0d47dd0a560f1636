// msda_fast.cuh -- vectorised, warp-cooperative MSDeformAttn kernels (sm_100a).
//
// Work decomposition (forward and backward):
//   * a GROUP of G lanes owns one (batch b, query q, head m); lane j of the
//     group owns NU 16-byte units of the head row (unit u = elements
//     [u*G*E + j*E, +E), E = 16/sizeof(T)) -> every tap is NU coalesced
//     128-bit loads per lane, each covering G*16 contiguous bytes.
//   * a CTA (256 threads = 256/G groups) owns a TILE = (b, m, chunk of
//     consecutive queries), so the rows tapped by neighbouring queries are
//     re-served from the SM's L1 instead of L2.
//   * the scalar geometry of a sampling point (floor, weights, gate) is
//     computed ONCE, by lane j for point j of the current chunk of G points,
//     and published to the group through shared memory (broadcast LDS.128)
//     -- the reference recomputes it in each of the D channel threads
//     (ms_deform_im2col_cuda.cuh:282-297).
//   * when every level has H == 1 (the temporal use of the op in this
//     reference, models/ops/modules/ms_deform_attn.py:114-117) only the two
//     taps of the single row are ever live; the kernels branch (uniformly)
//     to a 2-tap body.
#pragma once
#include "msda_common.cuh"
#include "msda_det.cuh"

namespace msda {

constexpr int kFastThreads = 256;

template <int NT> struct TapMeta;            // what a group needs per point
template <> struct __align__(16) TapMeta<2> { int off[2]; float w[2]; };
template <> struct __align__(16) TapMeta<4> { int off[4]; float w[4]; };

template <int G> __device__ __forceinline__ unsigned group_mask(int lane) {
  if constexpr (G == 32) {
    return 0xffffffffu;
  } else {
    return ((1u << G) - 1u) << (lane & ~(G - 1));
  }
}

// x[i] holds this lane's partial for point i (i < G).  On return lane j holds
// the sum over the group's lanes of x[j].  G-1 shuffles instead of G*log2(G).
template <int G>
__device__ __forceinline__ float transpose_reduce(float (&x)[G], int j, unsigned mask) {
#pragma unroll
  for (int s = G / 2; s >= 1; s >>= 1) {
    const bool up = (j & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? x[i] : x[i + s];
      const float keep = up ? x[i + s] : x[i];
      x[i] = keep + __shfl_xor_sync(mask, send, s);
    }
  }
  return x[0];
}

// Stage (H, W, start) per level and the "all H == 1" flag.
__device__ __forceinline__ void stage_levels(const int64_t* __restrict__ shapes,
                                             const int64_t* __restrict__ lsi, int L, int* s_lvl,
                                             int* s_one_d) {
  if (threadIdx.x == 0) *s_one_d = 1;
  __syncthreads();
  for (int l = threadIdx.x; l < L; l += blockDim.x) {
    const int H = (int)shapes[2 * l], W = (int)shapes[2 * l + 1];
    s_lvl[3 * l] = H;
    s_lvl[3 * l + 1] = W;
    s_lvl[3 * l + 2] = (int)lsi[l];
    if (H != 1) *s_one_d = 0;
  }
  __syncthreads();
}

template <typename T> struct LocLoad;
template <> struct LocLoad<float> {
  static __device__ __forceinline__ void ld(const float* loc2, const float* aw, float& x, float& y,
                                            float& a) {
    const float2 t = __ldcs(reinterpret_cast<const float2*>(loc2));
    x = t.x; y = t.y; a = __ldcs(aw);
  }
  static __device__ __forceinline__ void st(float* gloc2, float* gaw, float gx, float gy, float ga) {
    __stcs(reinterpret_cast<float2*>(gloc2), make_float2(gx, gy));
    __stcs(gaw, ga);
  }
};
template <> struct LocLoad<__nv_bfloat16> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* loc2, const __nv_bfloat16* aw,
                                            float& x, float& y, float& a) {
    const uint32_t t = __ldcs(reinterpret_cast<const unsigned int*>(loc2));
    x = __uint_as_float(t << 16); y = __uint_as_float(t & 0xffff0000u);
    a = __bfloat162float(*aw);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* gloc2, __nv_bfloat16* gaw, float gx,
                                            float gy, float ga) {
    *reinterpret_cast<__nv_bfloat162*>(gloc2) = __floats2bfloat162_rn(gx, gy);
    *gaw = __float2bfloat16_rn(ga);
  }
};

// ---- predicated 128-bit accesses (branch-free inner loops) ------------------
// E elements of T (16 bytes) -> E floats; all zeros when !pred (nothing is read).
template <typename T> struct Unit;
template <> struct Unit<float> {
  static constexpr int E = 4;
  static constexpr int RW = 4;      // raw 32-bit words of a unit as loaded
  static constexpr int BYTES = 16;
  static __device__ __forceinline__ void ldg(const float* p, bool pred, float (&v)[4]) {
    asm("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
        "mov.f32 %2, 0f00000000;\n\tmov.f32 %3, 0f00000000;\n\t"
        "@p ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];\n\t}"
        : "=&f"(v[0]), "=&f"(v[1]), "=&f"(v[2]), "=&f"(v[3])
        : "l"(p), "r"((int)pred));
  }
  using Raw = float;
  // predicated 128-bit load that KEEPS the old register contents when !pred
  static __device__ __forceinline__ void ldg_keep(const char* p, bool pred, float (&v)[4]) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3])
                 : "l"(p), "r"((int)pred));
  }
  // the same for data read once (grad_out): do not displace value rows from L1
  static __device__ __forceinline__ void ldg_keep_stream(const char* p, bool pred, float (&v)[4]) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3])
                 : "l"(p), "r"((int)pred));
  }
  static __device__ __forceinline__ void ldg_raw(const char* p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void unpack_raw(const float (&r)[4], float (&v)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = r[i];
  }
  static __device__ __forceinline__ void ldg_all(const float* p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void ld_stream(const float* p, float (&v)[4]) {
    const float4 t = __ldcs(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void st_stream(float* p, const float (&v)[4]) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
  }
};
template <> struct Unit<__nv_bfloat16> {
  static constexpr int E = 8;
  static constexpr int RW = 4;
  static constexpr int BYTES = 16;
  static __device__ __forceinline__ void unpack(const uint32_t (&w)[4], float (&v)[8]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void ldg(const __nv_bfloat16* p, bool pred, float (&v)[8]) {
    uint32_t w[4];
    asm("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
        "mov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
        "@p ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];\n\t}"
        : "=&r"(w[0]), "=&r"(w[1]), "=&r"(w[2]), "=&r"(w[3])
        : "l"(p), "r"((int)pred));
    unpack(w, v);
  }
  using Raw = uint32_t;
  static __device__ __forceinline__ void ldg_keep(const char* p, bool pred, uint32_t (&w)[4]) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3])
                 : "l"(p), "r"((int)pred));
  }
  static __device__ __forceinline__ void ldg_keep_stream(const char* p, bool pred, uint32_t (&w)[4]) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                 "@p ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];\n\t}"
                 : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3])
                 : "l"(p), "r"((int)pred));
  }
  static __device__ __forceinline__ void ldg_raw(const char* p, uint32_t (&w)[4]) {
    const uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
    w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
  }
  static __device__ __forceinline__ void unpack_raw(const uint32_t (&w)[4], float (&v)[8]) {
    unpack(w, v);
  }
  static __device__ __forceinline__ void ldg_all(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
    unpack(w, v);
  }
  static __device__ __forceinline__ void ld_stream(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 t = __ldcs(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
    unpack(w, v);
  }
  static __device__ __forceinline__ void st_stream(__nv_bfloat16* p, const float (&v)[8]) {
    VecIO<__nv_bfloat16, 8>::st_stream(p, v);
  }
};

// 8-byte units of 4 bf16, for the backward kernels: with two of them per lane (NU = 2, 32
// channels apart) a lane's slice of the fp32 accumulation image is two 16-byte pieces 128 bytes
// apart, so each 128-bit red of an 8-lane group covers a dense 128 bytes.  (With the 16-byte
// unit a lane owns 32 contiguous bytes of the image and every red instruction half-fills eight
// 32-byte sectors: twice the L2 atomic work for the same bytes.)
struct UnitBf16x4 {
  static constexpr int E = 4;
  static constexpr int RW = 2;
  static constexpr int BYTES = 8;
  using Raw = uint32_t;
  static __device__ __forceinline__ void unpack(const uint32_t (&w)[2], float (&v)[4]) {
    v[0] = __uint_as_float(w[0] << 16); v[1] = __uint_as_float(w[0] & 0xffff0000u);
    v[2] = __uint_as_float(w[1] << 16); v[3] = __uint_as_float(w[1] & 0xffff0000u);
  }
  static __device__ __forceinline__ void ldg(const __nv_bfloat16* p, bool pred, float (&v)[4]) {
    uint32_t w[2];
    asm("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
        "mov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\t"
        "@p ld.global.nc.v2.b32 {%0, %1}, [%2];\n\t}"
        : "=&r"(w[0]), "=&r"(w[1])
        : "l"(p), "r"((int)pred));
    unpack(w, v);
  }
  static __device__ __forceinline__ void ldg_keep(const char* p, bool pred, uint32_t (&w)[2]) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
                 "@p ld.global.nc.v2.b32 {%0, %1}, [%2];\n\t}"
                 : "+r"(w[0]), "+r"(w[1])
                 : "l"(p), "r"((int)pred));
  }
  static __device__ __forceinline__ void ldg_keep_stream(const char* p, bool pred, uint32_t (&w)[2]) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t"
                 "@p ld.global.nc.L1::no_allocate.v2.b32 {%0, %1}, [%2];\n\t}"
                 : "+r"(w[0]), "+r"(w[1])
                 : "l"(p), "r"((int)pred));
  }
  static __device__ __forceinline__ void ldg_raw(const char* p, uint32_t (&w)[2]) {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
    w[0] = t.x; w[1] = t.y;
  }
  static __device__ __forceinline__ void unpack_raw(const uint32_t (&w)[2], float (&v)[4]) { unpack(w, v); }
  static __device__ __forceinline__ void ld_stream(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 t = __ldcs(reinterpret_cast<const uint2*>(p));
    const uint32_t w[2] = {t.x, t.y};
    unpack(w, v);
  }
};

// predicated fire-and-forget 128-bit fp32 reduction (sm_90+ `red.global.add.v4.f32`)
__device__ __forceinline__ void red_add_f32x4_pred(float* p, bool pred, float a, float b, float c,
                                                   float d) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %5, 0;\n\t"
               "@p red.global.add.v4.f32 [%0], {%1, %2, %3, %4};\n\t}" ::"l"(p),
               "f"(a), "f"(b), "f"(c), "f"(d), "r"((int)pred)
               : "memory");
}

// grad_value accumulation of the vectorised kernels: fp32 reds (default), or 64-bit fixed
// point with integer atomics for the deterministic variant (msda_det.cuh)
template <typename GV> struct GvRed;
template <> struct GvRed<float> {
  static __device__ __forceinline__ void add4(float* p, bool pred, float a, float b, float c, float d,
                                              float) {
    red_add_f32x4_pred(p, pred, a, b, c, d);
  }
};
template <> struct GvRed<long long> {
  static __device__ __forceinline__ void add4(long long* p, bool pred, float a, float b, float c,
                                              float d, float scale) {
    if (pred) {
      unsigned long long* q = reinterpret_cast<unsigned long long*>(p);
      atomicAdd(q + 0, (unsigned long long)__float2ll_rn(a * scale));
      atomicAdd(q + 1, (unsigned long long)__float2ll_rn(b * scale));
      atomicAdd(q + 2, (unsigned long long)__float2ll_rn(c * scale));
      atomicAdd(q + 3, (unsigned long long)__float2ll_rn(d * scale));
    }
  }
};

// Which taps a 2-tap (H == 1) body uses: the top pair (0,1) when row h_low is
// the map's single row, else the bottom pair (2,3).
template <int NT> __device__ __forceinline__ bool use_bottom_pair(const PointGeom<float>& g) {
  if (NT == 4) return false;
  return !(g.off[0] >= 0 || g.off[1] >= 0);
}

// Owner lane: geometry of point `pt` of group `grp` -> TapMeta (element offsets
// row*M*D, weights pre-multiplied by the attention weight, invalid taps and
// points beyond LP neutralised) and, for the backward, the coefficients that
// turn the reduced dot products into the three gradients.
template <typename T, int NT, bool BWD>
__device__ __forceinline__ void owner_point(const T* __restrict__ loc_g, const T* __restrict__ aw_g,
                                            int pt, int LP, int P, int MD, bool border,
                                            const int* s_lvl, TapMeta<NT>& tm, float (&cw)[NT],
                                            float (&cx)[NT], float (&cy)[NT],
                                            const float* lpts = nullptr,
                                            const unsigned char* mrow = nullptr) {
#pragma unroll
  for (int k = 0; k < NT; ++k) { tm.off[k] = -1; tm.w[k] = 0.f; cw[k] = cx[k] = cy[k] = 0.f; }
  if (pt >= LP) return;
  float lx, ly, a;
  if (lpts) {   // (x, y, weight) of the query's points as floats, e.g. in shared memory (fused front-end)
    lx = lpts[3 * pt]; ly = lpts[3 * pt + 1]; a = lpts[3 * pt + 2];
  } else {
    LocLoad<T>::ld(loc_g + 2 * pt, aw_g + pt, lx, ly, a);
  }
  const int l = pt / P;
  PointGeom<float> g;
  point_geom<float>(lx, ly, s_lvl[3 * l], s_lvl[3 * l + 1], s_lvl[3 * l + 2], border, g);
  const bool alt = use_bottom_pair<NT>(g);
  const float hh = 1.f - g.lh, hw = 1.f - g.lw;
  const float dx[4] = {-hh, hh, -g.lh, g.lh};
  const float dy[4] = {-hw, -g.lw, hw, g.lw};
#pragma unroll
  for (int k = 0; k < NT; ++k) {
    const int k2 = (k + 2) & 3;
    const int row = alt ? g.off[k2] : g.off[k];
    const float wk = alt ? g.w[k2] : g.w[k];
    // mrow: padding mask of the batch item (fused module op): a masked row counts as zeros, i.e. is
    // neither read nor given a gradient (value.masked_fill(mask, 0), ms_deform_attn.py:96-97)
    const bool ok = row >= 0 && !(mrow && mrow[row >= 0 ? row : 0]);
    tm.off[k] = ok ? row * MD : -1;
    tm.w[k] = ok ? wk * a : 0.f;
    if (BWD) {
      cw[k] = wk;
      cx[k] = (alt ? dx[k2] : dx[k]) * g.sx * a;
      cy[k] = (alt ? dy[k2] : dy[k]) * g.sy * a;
    }
  }
}

// ------------------------------- forward ------------------------------------
template <typename T, int NU, int G, int NT>
__device__ __forceinline__ void fwd_body(const T* __restrict__ value, const T* __restrict__ loc,
                                         const T* __restrict__ aw, T* __restrict__ out,
                                         const Dims& d, int b, int m, int q_first, int q_end,
                                         int q_step, const int* s_lvl, TapMeta<NT>* s_meta,
                                         const int* qlist = nullptr, const float* lpts = nullptr,
                                         const unsigned char* mrow = nullptr) {
  // the group handles queries q_first, q_first + q_step, ... < q_end of (b, m) -- or, with a
  // query list (the time-sorted tiles of msda_tile.cuh), the queries qlist[q_first], ...,
  // negative entries skipped; s_meta = G entries private to the group
  constexpr int E = Unit<T>::E;
  const int lane = threadIdx.x & 31;
  const int j = threadIdx.x % G;
  const unsigned mask = group_mask<G>(lane);
  const int LP = d.L * d.P;
  const int MD = d.M * d.D;
  const T* vb = value + ((size_t)b * d.S * d.M + m) * d.D + j * E;

  for (int qq = q_first; qq < q_end; qq += q_step) {
    const int q = qlist ? qlist[qq] : qq;
    if (q < 0) continue;
    const size_t grp = ((size_t)b * d.Lq + q) * d.M + m;
    const T* loc_g = loc + grp * LP * 2;
    const T* aw_g = aw + grp * LP;
    float2 acc[NU][E / 2];
#pragma unroll
    for (int u = 0; u < NU; ++u)
#pragma unroll
      for (int c = 0; c < E / 2; ++c) acc[u][c] = make_float2(0.f, 0.f);

    for (int base = 0; base < LP; base += G) {
      {
        TapMeta<NT> tm;
        float cw[NT], cx[NT], cy[NT];
        owner_point<T, NT, false>(loc_g, aw_g, base + j, LP, d.P, MD, d.border != 0, s_lvl, tm, cw,
                                  cx, cy, lpts ? lpts + (size_t)qq * LP * 3 : nullptr, mrow);
        s_meta[j] = tm;
      }
      __syncwarp(mask);
#pragma unroll
      for (int i = 0; i < G; ++i) {
        const TapMeta<NT> tm = s_meta[i];
#pragma unroll
        for (int k = 0; k < NT; ++k) {
          const float2 w2 = make_float2(tm.w[k], tm.w[k]);
#pragma unroll
          for (int u = 0; u < NU; ++u) {
            float v[E];
            Unit<T>::ldg(vb + tm.off[k] + u * (G * E), tm.off[k] >= 0, v);
#pragma unroll
            for (int c = 0; c < E / 2; ++c)
              acc[u][c] = __ffma2_rn(w2, make_float2(v[2 * c], v[2 * c + 1]), acc[u][c]);
          }
        }
      }
      __syncwarp(mask);
    }
#pragma unroll
    for (int u = 0; u < NU; ++u) {
      float o[E];
#pragma unroll
      for (int c = 0; c < E / 2; ++c) { o[2 * c] = acc[u][c].x; o[2 * c + 1] = acc[u][c].y; }
      Unit<T>::st_stream(out + grp * d.D + u * (G * E) + j * E, o);
    }
  }
}

template <typename T, int NU, int G>
__global__ void __launch_bounds__(kFastThreads)
msda_fwd_fast_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                     const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                     const T* __restrict__ aw, T* __restrict__ out, Dims d, int qtile, int nqc) {
  __shared__ int s_lvl[3 * kMaxLevelsFast];
  __shared__ int s_one_d;
  __shared__ TapMeta<4> s_meta[kFastThreads];
  stage_levels(shapes, lsi, d.L, s_lvl, &s_one_d);

  const int tile = blockIdx.x;
  const int m = tile % d.M;
  const int qc = (tile / d.M) % nqc;
  const int b = tile / (d.M * nqc);
  const int q_begin = qc * qtile;
  const int q_end = min(q_begin + qtile, d.Lq);
  constexpr int NG = kFastThreads / G;
  const int gi = threadIdx.x / G;
  if (s_one_d)
    fwd_body<T, NU, G, 2>(value, loc, aw, out, d, b, m, q_begin + gi, q_end, NG, s_lvl,
                          reinterpret_cast<TapMeta<2>*>(s_meta) + gi * G);
  else
    fwd_body<T, NU, G, 4>(value, loc, aw, out, d, b, m, q_begin + gi, q_end, NG, s_lvl,
                          s_meta + gi * G);
}

// ------------------------------- backward -----------------------------------
// grad_value is accumulated in fp32 with 128-bit `red.global.add.v4.f32`
// (for bf16 storage into an fp32 workspace image that a second kernel rounds).
// grad_sampling_loc / grad_attn_weight need, per point, the dot products
// <grad_out, v_k> over the head's D channels: each lane keeps its partial for
// the G points of the chunk and one transpose-reduce hands point j's totals to
// lane j, which owns that point's geometry and writes its three gradients
// (.cuh:115-159 for the math).
template <typename T, int NU, int G, int NT, typename GV = float, typename U = Unit<T>>
__device__ __forceinline__ void bwd_body(const T* __restrict__ value, const T* __restrict__ loc,
                                         const T* __restrict__ aw, const T* __restrict__ gout,
                                         GV* __restrict__ gvalue, T* __restrict__ gloc,
                                         T* __restrict__ gaw, const Dims& d, int b, int m,
                                         int q_first, int q_end, int q_step, const int* s_lvl,
                                         TapMeta<NT>* s_meta, float gv_scale = 1.f,
                                         const int* qlist = nullptr, const float* lpts = nullptr,
                                         float* lgrad = nullptr, const unsigned char* mrow = nullptr) {
  constexpr int E = U::E;
  const int lane = threadIdx.x & 31;
  const int j = threadIdx.x % G;
  const unsigned mask = group_mask<G>(lane);
  const int LP = d.L * d.P;
  const int MD = d.M * d.D;
  const size_t voff = ((size_t)b * d.S * d.M + m) * d.D + j * E;
  const T* vb = value + voff;
  GV* gvb = gvalue + voff;

  for (int qq = q_first; qq < q_end; qq += q_step) {
    const int q = qlist ? qlist[qq] : qq;
    if (q < 0) continue;
    const size_t grp = ((size_t)b * d.Lq + q) * d.M + m;
    const T* loc_g = loc + grp * LP * 2;
    const T* aw_g = aw + grp * LP;
    float go[NU][E];
#pragma unroll
    for (int u = 0; u < NU; ++u) U::ld_stream(gout + grp * d.D + u * (G * E) + j * E, go[u]);

    for (int base = 0; base < LP; base += G) {
      const int pt = base + j;
      float cw[NT], cx[NT], cy[NT];
      {
        TapMeta<NT> tm;
        owner_point<T, NT, true>(loc_g, aw_g, pt, LP, d.P, MD, d.border != 0, s_lvl, tm, cw, cx, cy,
                                 lpts ? lpts + (size_t)qq * LP * 3 : nullptr, mrow);
        s_meta[j] = tm;
      }
      __syncwarp(mask);
      float dot[NT][G];
#pragma unroll
      for (int i = 0; i < G; ++i) {
        const TapMeta<NT> tm = s_meta[i];
#pragma unroll
        for (int k = 0; k < NT; ++k) {
          const bool live = tm.off[k] >= 0;
          float s = 0.f;
#pragma unroll
          for (int u = 0; u < NU; ++u) {
            float v[E];
            U::ldg(vb + tm.off[k] + u * (G * E), live, v);
#pragma unroll
            for (int c = 0; c < E; ++c) s = fmaf(go[u][c], v[c], s);
#pragma unroll
            for (int c = 0; c < E; c += 4)
              GvRed<GV>::add4(gvb + tm.off[k] + u * (G * E) + c, live, tm.w[k] * go[u][c],
                              tm.w[k] * go[u][c + 1], tm.w[k] * go[u][c + 2],
                              tm.w[k] * go[u][c + 3], gv_scale);
          }
          dot[k][i] = s;
        }
      }
      float ga = 0.f, gx = 0.f, gy = 0.f;
#pragma unroll
      for (int k = 0; k < NT; ++k) {
        const float r = transpose_reduce<G>(dot[k], j, mask);
        ga = fmaf(cw[k], r, ga);
        gx = fmaf(cx[k], r, gx);
        gy = fmaf(cy[k], r, gy);
      }
      if (pt < LP) {
        if (lgrad) {   // (d/dx, d/dy, d/d weight) per point as floats, for a fused epilogue
          float* gq = lgrad + ((size_t)qq * LP + pt) * 3;
          gq[0] = gx; gq[1] = gy; gq[2] = ga;
        } else {
          LocLoad<T>::st(gloc + (grp * LP + pt) * 2, gaw + grp * LP + pt, gx, gy, ga);
        }
      }
      __syncwarp(mask);
    }
  }
}

template <typename T, int NU, int G, typename GV = float, typename U = Unit<T>>
__global__ void __launch_bounds__(kFastThreads)
msda_bwd_fast_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                     const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                     const T* __restrict__ aw, const T* __restrict__ gout,
                     GV* __restrict__ gvalue, T* __restrict__ gloc, T* __restrict__ gaw, Dims d,
                     int qtile, int nqc, const DetHeader* __restrict__ det) {
  const float gv_scale = det ? (float)det->scale : 1.f;
  __shared__ int s_lvl[3 * kMaxLevelsFast];
  __shared__ int s_one_d;
  __shared__ TapMeta<4> s_meta[kFastThreads];
  stage_levels(shapes, lsi, d.L, s_lvl, &s_one_d);

  const int tile = blockIdx.x;
  const int m = tile % d.M;
  const int qc = (tile / d.M) % nqc;
  const int b = tile / (d.M * nqc);
  const int q_begin = qc * qtile;
  const int q_end = min(q_begin + qtile, d.Lq);
  constexpr int NG = kFastThreads / G;
  const int gi = threadIdx.x / G;
  if (s_one_d)
    bwd_body<T, NU, G, 2, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, q_begin + gi, q_end,
                              NG, s_lvl, reinterpret_cast<TapMeta<2>*>(s_meta) + gi * G, gv_scale);
  else
    bwd_body<T, NU, G, 4, GV, U>(value, loc, aw, gout, gvalue, gloc, gaw, d, b, m, q_begin + gi, q_end,
                              NG, s_lvl, s_meta + gi * G, gv_scale);
}

// fp32 accumulation image -> bf16 grad_value
__global__ void __launch_bounds__(256)
msda_round_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8;
       i += (size_t)gridDim.x * blockDim.x) {
    const float4 a = __ldcs(reinterpret_cast<const float4*>(src) + 2 * i);
    const float4 b = __ldcs(reinterpret_cast<const float4*>(src) + 2 * i + 1);
    const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    VecIO<__nv_bfloat16, 8>::st_stream(dst + 8 * i, v);
  }
}

}  // namespace msda
