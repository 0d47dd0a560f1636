// msda_generic.cuh -- any-D / any-dtype (f32, f64, bf16) MSDeformAttn kernels.
// Correctness-first path used when the vectorised kernels do not apply
// (D*sizeof(T) not a multiple of 16 B, D/VEC not a power of two <= 32, fp64,
// more than kMaxLevelsFast levels).  Covers the odd channel counts the
// reference tests with gradcheck (models/ops/test.py:85: 30, 71, 1025, 3096).
#pragma once
#include "msda_common.cuh"
#include "msda_det.cuh"

namespace msda {

// one thread per output element, grid-stride (cf. .cuh:238-300)
template <typename T>
__global__ void __launch_bounds__(256)
msda_fwd_generic_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                        const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                        const T* __restrict__ aw, T* __restrict__ out, Dims d, size_t total) {
  using CT = typename AccOf<T>::type;
  const size_t MD = (size_t)d.M * d.D;
  const int LP = d.L * d.P;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % d.D);
    const size_t grp = idx / d.D;  // (b*Lq + q)*M + m
    const int m = (int)(grp % d.M);
    const size_t b = grp / ((size_t)d.M * d.Lq);
    const T* vb = value + b * d.S * MD + (size_t)m * d.D + c;
    const T* loc_g = loc + grp * LP * 2;
    const T* aw_g = aw + grp * LP;
    CT acc = (CT)0;
    for (int l = 0; l < d.L; ++l) {
      const int H = (int)shapes[2 * l], W = (int)shapes[2 * l + 1], start = (int)lsi[l];
      for (int p = 0; p < d.P; ++p) {
        const int pt = l * d.P + p;
        PointGeom<CT> g;
        point_geom<CT>(to_acc(loc_g[2 * pt]), to_acc(loc_g[2 * pt + 1]), H, W, start,
                       d.border != 0, g);
        CT val = (CT)0;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (g.off[k] >= 0) val += g.w[k] * to_acc(vb[(size_t)g.off[k] * MD]);
        acc += val * to_acc(aw_g[pt]);
      }
    }
    out[idx] = from_acc<T, CT>(acc);
  }
}

template <typename CT> __device__ __forceinline__ CT warp_sum(CT v) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
  return v;
}

// one CTA per (b, q, m); threads stride over the D channels.  GV is the
// grad_value accumulation type (T for f32/f64, float workspace for bf16, int64
// fixed point with `det` for the deterministic variant).
template <typename T, typename GV>
__global__ void __launch_bounds__(128)
msda_bwd_generic_kernel(const T* __restrict__ value, const int64_t* __restrict__ shapes,
                        const int64_t* __restrict__ lsi, const T* __restrict__ loc,
                        const T* __restrict__ aw, const T* __restrict__ gout,
                        GV* __restrict__ gvalue, T* __restrict__ gloc, T* __restrict__ gaw,
                        Dims d, const DetHeader* __restrict__ det) {
  using CT = typename AccOf<T>::type;
  __shared__ CT s_part[4][4];  // [warp][tap]
  const CT gv_scale = det ? (CT)det->scale : (CT)1;   // fixed-point accumulation (msda_det.cuh)
  const size_t grp = blockIdx.x;
  const int m = (int)(grp % d.M);
  const size_t b = grp / ((size_t)d.M * d.Lq);
  const size_t MD = (size_t)d.M * d.D;
  const int LP = d.L * d.P;
  const size_t voff = b * d.S * MD + (size_t)m * d.D;
  const T* loc_g = loc + grp * LP * 2;
  const T* aw_g = aw + grp * LP;
  const T* go_g = gout + grp * d.D;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int l = 0; l < d.L; ++l) {
    const int H = (int)shapes[2 * l], W = (int)shapes[2 * l + 1], start = (int)lsi[l];
    for (int p = 0; p < d.P; ++p) {
      const int pt = l * d.P + p;
      const CT a = to_acc(aw_g[pt]);
      PointGeom<CT> g;
      point_geom<CT>(to_acc(loc_g[2 * pt]), to_acc(loc_g[2 * pt + 1]), H, W, start, d.border != 0,
                     g);
      CT dot[4] = {(CT)0, (CT)0, (CT)0, (CT)0};
      for (int c = threadIdx.x; c < d.D; c += blockDim.x) {
        const CT gc = to_acc(go_g[c]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (g.off[k] >= 0) {
            const size_t o = voff + (size_t)g.off[k] * MD + c;
            dot[k] += gc * to_acc(value[o]);
            AccAdd<GV, CT>::add(gvalue + o, g.w[k] * a * gc, gv_scale);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        dot[k] = warp_sum(dot[k]);
        if (lane == 0) s_part[warp][k] = dot[k];
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        CT r[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          r[k] = s_part[0][k] + s_part[1][k] + s_part[2][k] + s_part[3][k];
        const CT hh = (CT)1 - g.lh, hw = (CT)1 - g.lw;
        const CT ga = g.w[0] * r[0] + g.w[1] * r[1] + g.w[2] * r[2] + g.w[3] * r[3];
        const CT gx = g.sx * a * (-hh * r[0] + hh * r[1] - g.lh * r[2] + g.lh * r[3]);
        const CT gy = g.sy * a * (-hw * r[0] - g.lw * r[1] + hw * r[2] + g.lw * r[3]);
        gaw[grp * LP + pt] = from_acc<T, CT>(ga);
        gloc[(grp * LP + pt) * 2] = from_acc<T, CT>(gx);
        gloc[(grp * LP + pt) * 2 + 1] = from_acc<T, CT>(gy);
      }
      __syncthreads();
    }
  }
}

// fp32 accumulation image -> bf16, scalar tail-safe version
__global__ void __launch_bounds__(256)
msda_round_bf16_scalar_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                              size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    dst[i] = __float2bfloat16_rn(src[i]);
}

}  // namespace msda
