// msda_common.cuh -- shared device helpers for the B200 MSDeformAttn kernels.
//
// Semantics follow the reference op (models/ops/src/cuda/ms_deform_im2col_cuda.cuh):
//   pixel coords  w_im = loc_x*W - 0.5, h_im = loc_y*H - 0.5            (:286-287)
//   zeros gate    h_im>-1 && w_im>-1 && h_im<H && w_im<W                (:289)
//   taps          low=floor, high=low+1, weights hh*hw, hh*lw, lh*hw, lh*lw,
//                 a tap is read only when inside the map                (:39-81)
// plus a `border` mode that reproduces F.grid_sample(padding_mode='border'),
// which is what the reference's PyTorch path uses
// (models/ops/functions/ms_deform_attn_func.py:61-62).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace msda {

// accumulation type: fp32 for f32/bf16 storage, fp64 for f64 storage
template <typename T> struct AccOf { using type = float; };
template <> struct AccOf<double> { using type = double; };

template <typename T> __device__ __forceinline__ typename AccOf<T>::type to_acc(T v) { return v; }
template <> __device__ __forceinline__ float to_acc<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}
template <typename T, typename CT> __device__ __forceinline__ T from_acc(CT v) { return (T)v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16, float>(float v) {
  return __float2bfloat16_rn(v);
}

template <typename CT> __device__ __forceinline__ CT floor_t(CT x);
template <> __device__ __forceinline__ float floor_t<float>(float x) { return floorf(x); }
template <> __device__ __forceinline__ double floor_t<double>(double x) { return floor(x); }

// Geometry of one sampling point: which rows are tapped and with which weights.
// off[k] is the row index inside one batch item (level_start + h*W + w), or -1
// when tap k is not read.  Tap order = reference v1..v4:
//   0:(h_low,w_low) 1:(h_low,w_high) 2:(h_high,w_low) 3:(h_high,w_high)
template <typename CT>
struct PointGeom {
  int off[4];
  CT w[4];    // hh*hw, hh*lw, lh*hw, lh*lw
  CT lh, lw;  // fractional parts (hh = 1-lh, hw = 1-lw)
  CT sx, sy;  // d(w_im)/d(loc_x), d(h_im)/d(loc_y): W,H or 0 (gated / clamped)
};

template <typename CT>
__device__ __forceinline__ void point_geom(CT loc_x, CT loc_y, int H, int W, int start,
                                           bool border, PointGeom<CT>& g) {
  CT h_im = loc_y * (CT)H - (CT)0.5;
  CT w_im = loc_x * (CT)W - (CT)0.5;
  CT sx = (CT)W, sy = (CT)H;
  bool live;
  if (border) {
    // grid_sample clip_coordinates_set_grad: the borders themselves count as
    // out of bounds for the gradient.
    const CT wmax = (CT)(W - 1), hmax = (CT)(H - 1);
    if (!(w_im > (CT)0)) { w_im = (CT)0; sx = (CT)0; }
    else if (w_im >= wmax) { w_im = wmax; sx = (CT)0; }
    if (!(h_im > (CT)0)) { h_im = (CT)0; sy = (CT)0; }
    else if (h_im >= hmax) { h_im = hmax; sy = (CT)0; }
    live = true;
  } else {
    live = (h_im > (CT)-1) && (w_im > (CT)-1) && (h_im < (CT)H) && (w_im < (CT)W);
  }
  if (!live) {
#pragma unroll
    for (int k = 0; k < 4; ++k) { g.off[k] = -1; g.w[k] = (CT)0; }
    g.lh = g.lw = (CT)0;
    g.sx = g.sy = (CT)0;
    return;
  }
  const CT hf = floor_t(h_im), wf = floor_t(w_im);
  const int h_low = (int)hf, w_low = (int)wf;
  const int h_high = h_low + 1, w_high = w_low + 1;
  const CT lh = h_im - hf, lw = w_im - wf;
  const CT hh = (CT)1 - lh, hw = (CT)1 - lw;
  const bool hl = h_low >= 0, hhi = h_high <= H - 1;
  const bool wl = w_low >= 0, whi = w_high <= W - 1;
  const int r0 = start + h_low * W + w_low;
  g.off[0] = (hl && wl) ? r0 : -1;
  g.off[1] = (hl && whi) ? r0 + 1 : -1;
  g.off[2] = (hhi && wl) ? r0 + W : -1;
  g.off[3] = (hhi && whi) ? r0 + W + 1 : -1;
  g.w[0] = hh * hw; g.w[1] = hh * lw; g.w[2] = lh * hw; g.w[3] = lh * lw;
  g.lh = lh; g.lw = lw;
  g.sx = sx; g.sy = sy;
}

// ---- vector global memory access -------------------------------------------
// VEC elements of T starting at p (16-byte aligned when VEC*sizeof(T)==16).
template <typename T, int VEC> struct VecIO;

template <> struct VecIO<float, 4> {
  static __device__ __forceinline__ void ldg(const float* p, float (&v)[4]) {
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

template <> struct VecIO<__nv_bfloat16, 8> {
  static __device__ __forceinline__ void unpack(const uint4& t, float (&v)[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void ldg(const __nv_bfloat16* p, float (&v)[8]) {
    unpack(__ldg(reinterpret_cast<const uint4*>(p)), v);
  }
  static __device__ __forceinline__ void ld_stream(const __nv_bfloat16* p, float (&v)[8]) {
    unpack(__ldcs(reinterpret_cast<const uint4*>(p)), v);
  }
  static __device__ __forceinline__ void st_stream(__nv_bfloat16* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    __stcs(reinterpret_cast<uint4*>(p), make_uint4(w[0], w[1], w[2], w[3]));
  }
};

// fire-and-forget 128-bit fp32 reduction into global memory (sm_90+ `red.v4.f32`)
__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}

// Level table staged in shared memory: (H, W, start) per level.
constexpr int kMaxLevelsFast = 32;

struct Dims {
  int N, S, M, D, L, Lq, P;
  int border;
};

}  // namespace msda
