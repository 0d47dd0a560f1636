// msda_frontend.cuh -- the module's front-end in one kernel each way.
//
// MSDeformAttn.forward (models/ops/modules/ms_deform_attn.py:98-117; live twin
// models/modules/attention.py:462-496) turns the two Linear outputs of a query
// into the op's arguments with a softmax and four or five elementwise kernels:
//   aw  = softmax(logits over L*P)                                        (:100)
//   x   = ref[..,l,0] + off / T_l                       (ref has 1 column, :102-105)
//       = ref[..,l,0] + off / P * ref[..,l,1] * 0.5     (ref has 2 columns, :106-108)
//   loc = (x, 0.5)                                                        (:114-116)
// Here: one thread per (batch, query, head) reads its L*P offsets and logits once
// (16-byte pieces) and writes loc and aw; the backward takes (grad_loc, grad_aw),
// the saved aw and offsets, and writes grad_offsets, grad_logits and grad_ref
// (summed over heads and points with shuffles when M is a power of two, with
// atomics otherwise).
//
// Pure streaming: algorithmic bytes forward = e * groups * L*P * (2 read + 3 written),
// ENC fp32: 12.3 + 12.3 MB read, 24.6 + 12.3 MB written.
#pragma once
#include "msda_common.cuh"

namespace msda {

constexpr int kFrontThreads = 128;

// V consecutive elements of T as accumulation-type values (V = 4: one 8/16/32-byte piece)
template <typename T, int V> struct FrontVec;
template <typename T> struct FrontVec<T, 1> {
  using CT = typename AccOf<T>::type;
  static __device__ __forceinline__ void ld(const T* p, CT (&v)[1]) { v[0] = to_acc<T>(p[0]); }
  static __device__ __forceinline__ void st(T* p, const CT (&v)[1]) { p[0] = from_acc<T, CT>(v[0]); }
};
template <> struct FrontVec<float, 4> {
  static __device__ __forceinline__ void ld(const float* p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void st(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct FrontVec<double, 4> {
  static __device__ __forceinline__ void ld(const double* p, double (&v)[4]) {
    const double2 a = *reinterpret_cast<const double2*>(p);
    const double2 b = *(reinterpret_cast<const double2*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
  }
  static __device__ __forceinline__ void st(double* p, const double (&v)[4]) {
    *reinterpret_cast<double2*>(p) = make_double2(v[0], v[1]);
    *(reinterpret_cast<double2*>(p) + 1) = make_double2(v[2], v[3]);
  }
};
template <> struct FrontVec<__nv_bfloat16, 4> {
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 t = *reinterpret_cast<const uint2*>(p);
    v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
    v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, const float (&v)[4]) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<const uint32_t*>(&a);
    t.y = *reinterpret_cast<const uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = t;
  }
};

template <typename CT> __device__ __forceinline__ CT exp_t(CT x);
template <> __device__ __forceinline__ float exp_t<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ double exp_t<double>(double x) { return exp(x); }

struct FrontDims {
  long long groups;   // N * Lq * M
  int M, L, P;
  int ref_dim;        // 1: reference point, 2: (centre, length)
};

// x = ref0 + off * (1 / normaliser), written the way the module evaluates it
template <typename CT>
__device__ __forceinline__ CT front_x(CT off, CT r0, CT r1, CT t_l, CT p, int ref_dim) {
  return ref_dim == 1 ? r0 + off / t_l : r0 + off / p * r1 * (CT)0.5;
}

template <typename T, int V>
__global__ void __launch_bounds__(kFrontThreads)
msda_frontend_fwd_kernel(const T* __restrict__ off, const T* __restrict__ logits,
                         const T* __restrict__ ref, const int64_t* __restrict__ t_lvl,
                         T* __restrict__ loc, T* __restrict__ aw, FrontDims d) {
  using CT = typename AccOf<T>::type;
  const long long g = (long long)blockIdx.x * kFrontThreads + threadIdx.x;
  if (g >= d.groups) return;
  const int LP = d.L * d.P;
  const T* lg = logits + g * LP;
  const T* of = off + g * LP;
  const T* rf = ref + (g / d.M) * d.L * d.ref_dim;
  T* lo = loc + g * LP * 2;
  T* ao = aw + g * LP;

  CT mx = -INFINITY;
  for (int c = 0; c < LP; c += V) {
    CT v[V];
    FrontVec<T, V>::ld(lg + c, v);
#pragma unroll
    for (int k = 0; k < V; ++k) mx = v[k] > mx ? v[k] : mx;
  }
  CT sum = 0;
  for (int c = 0; c < LP; c += V) {
    CT v[V];
    FrontVec<T, V>::ld(lg + c, v);
#pragma unroll
    for (int k = 0; k < V; ++k) sum += exp_t<CT>(v[k] - mx);
  }
  const CT inv = (CT)1 / sum;
  for (int l = 0; l < d.L; ++l) {
    const CT r0 = to_acc<T>(rf[l * d.ref_dim]);
    const CT r1 = d.ref_dim == 2 ? to_acc<T>(rf[l * 2 + 1]) : (CT)0;
    const CT tl = (CT)t_lvl[l];
    for (int c = l * d.P; c < (l + 1) * d.P; c += V) {
      CT v[V], o[V];
      FrontVec<T, V>::ld(lg + c, v);
      FrontVec<T, V>::ld(of + c, o);
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = exp_t<CT>(v[k] - mx) * inv;
      FrontVec<T, V>::st(ao + c, v);
      if constexpr (V == 4) {
        // (x0, .5, x1, .5) (x2, .5, x3, .5)
        CT x0[4], x1[4];
        x0[0] = front_x<CT>(o[0], r0, r1, tl, (CT)d.P, d.ref_dim); x0[1] = (CT)0.5;
        x0[2] = front_x<CT>(o[1], r0, r1, tl, (CT)d.P, d.ref_dim); x0[3] = (CT)0.5;
        x1[0] = front_x<CT>(o[2], r0, r1, tl, (CT)d.P, d.ref_dim); x1[1] = (CT)0.5;
        x1[2] = front_x<CT>(o[3], r0, r1, tl, (CT)d.P, d.ref_dim); x1[3] = (CT)0.5;
        FrontVec<T, V>::st(lo + 2 * c, x0);
        FrontVec<T, V>::st(lo + 2 * c + V, x1);
      } else {
        lo[2 * c] = from_acc<T, CT>(front_x<CT>(o[0], r0, r1, tl, (CT)d.P, d.ref_dim));
        lo[2 * c + 1] = from_acc<T, CT>((CT)0.5);
      }
    }
  }
}

template <typename T> __device__ __forceinline__ void front_atomic_add(T* p, typename AccOf<T>::type v) {
  atomicAdd(p, (T)v);
}
template <> __device__ __forceinline__ void front_atomic_add<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  atomicAdd(p, __float2bfloat16_rn(v));
}

// SHFL: M is a power of two <= 32, so the M heads of a query sit in one warp and
// grad_ref is reduced with shuffles and stored once; otherwise it is accumulated
// with atomics into a zero-filled grad_ref.  grad_ref may be null (reference
// points that need no gradient, e.g. the encoder's).
template <typename T, int V, bool SHFL>
__global__ void __launch_bounds__(kFrontThreads)
msda_frontend_bwd_kernel(const T* __restrict__ off, const T* __restrict__ aw,
                         const T* __restrict__ ref, const int64_t* __restrict__ t_lvl,
                         const T* __restrict__ gloc, const T* __restrict__ gaw,
                         T* __restrict__ goff, T* __restrict__ glogits, T* __restrict__ gref,
                         FrontDims d) {
  using CT = typename AccOf<T>::type;
  const long long g = (long long)blockIdx.x * kFrontThreads + threadIdx.x;
  const bool valid = g < d.groups;
  const unsigned mask = __ballot_sync(0xffffffffu, valid);
  if (!valid) return;
  const int LP = d.L * d.P;
  const long long nq = g / d.M;
  const int m = (int)(g - nq * d.M);
  const T* awp = aw + g * LP;
  const T* gap = gaw + g * LP;
  const T* glp = gloc + g * LP * 2;
  const T* of = off + g * LP;
  const T* rf = ref + nq * d.L * d.ref_dim;

  CT dot = 0;
  for (int c = 0; c < LP; c += V) {
    CT a[V], ga[V];
    FrontVec<T, V>::ld(awp + c, a);
    FrontVec<T, V>::ld(gap + c, ga);
#pragma unroll
    for (int k = 0; k < V; ++k) dot += a[k] * ga[k];
  }
  for (int l = 0; l < d.L; ++l) {
    const CT r1 = d.ref_dim == 2 ? to_acc<T>(rf[l * 2 + 1]) : (CT)0;
    const CT tl = (CT)t_lvl[l];
    CT s0 = 0, s1 = 0;                  // sum_p gx ; sum_p gx * d x / d ref1
    for (int c = l * d.P; c < (l + 1) * d.P; c += V) {
      CT a[V], ga[V], gx[V], o[V];
      FrontVec<T, V>::ld(awp + c, a);
      FrontVec<T, V>::ld(gap + c, ga);
      FrontVec<T, V>::ld(of + c, o);
      if constexpr (V == 4) {
        CT t0[4], t1[4];
        FrontVec<T, V>::ld(glp + 2 * c, t0);
        FrontVec<T, V>::ld(glp + 2 * c + V, t1);
        gx[0] = t0[0]; gx[1] = t0[2]; gx[2] = t1[0]; gx[3] = t1[2];
      } else {
        gx[0] = to_acc<T>(glp[2 * c]);
      }
#pragma unroll
      for (int k = 0; k < V; ++k) {
        a[k] = a[k] * (ga[k] - dot);
        s0 += gx[k];
        s1 += gx[k] * (CT)0.5 * (o[k] / (CT)d.P);
        gx[k] = d.ref_dim == 1 ? gx[k] / tl : gx[k] * (CT)0.5 * r1 / (CT)d.P;
      }
      FrontVec<T, V>::st(glogits + g * LP + c, a);
      FrontVec<T, V>::st(goff + g * LP + c, gx);
    }
    if (gref != nullptr) {
      T* gr = gref + (nq * d.L + l) * d.ref_dim;
      if constexpr (SHFL) {
        for (int s = d.M >> 1; s > 0; s >>= 1) {
          s0 += __shfl_xor_sync(mask, s0, s);
          s1 += __shfl_xor_sync(mask, s1, s);
        }
        if (m == 0) {
          gr[0] = from_acc<T, CT>(s0);
          if (d.ref_dim == 2) gr[1] = from_acc<T, CT>(s1);
        }
      } else {
        front_atomic_add<T>(gr, s0);
        if (d.ref_dim == 2) front_atomic_add<T>(gr + 1, s1);
      }
    }
  }
}


// ---- lane-per-piece variants -------------------------------------------------
// When L*P/4 (= tpg, threads per group) is a power of two <= 32, thread i takes the
// i-th 4-element piece of the whole problem: every load and store is a fully
// coalesced 8/16/32-byte access and each array is touched exactly once; the
// softmax statistics travel over the tpg lanes of a group with shuffles, and
// grad_ref over the M heads of a query when those share a warp (M * tpg <= 32).
template <typename T>
__global__ void __launch_bounds__(kFrontThreads)
msda_frontend_fwd_lanes_kernel(const T* __restrict__ off, const T* __restrict__ logits,
                               const T* __restrict__ ref, const int64_t* __restrict__ t_lvl,
                               T* __restrict__ loc, T* __restrict__ aw, FrontDims d, int tpg) {
  using CT = typename AccOf<T>::type;
  const long long tid = (long long)blockIdx.x * kFrontThreads + threadIdx.x;
  const bool valid = tid < d.groups * tpg;
  const unsigned mask = __ballot_sync(0xffffffffu, valid);
  if (!valid) return;
  const long long g = tid / tpg;
  const int c = (int)(tid - g * tpg);             // piece inside the group
  const int l = c * 4 / d.P;
  CT v[4], o[4];
  FrontVec<T, 4>::ld(logits + tid * 4, v);
  FrontVec<T, 4>::ld(off + tid * 4, o);
  CT mx = fmax(fmax(v[0], v[1]), fmax(v[2], v[3]));
  for (int s = tpg >> 1; s > 0; s >>= 1) mx = fmax(mx, __shfl_xor_sync(mask, mx, s));
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = exp_t<CT>(v[k] - mx);
  CT sum = (v[0] + v[1]) + (v[2] + v[3]);
  for (int s = tpg >> 1; s > 0; s >>= 1) sum += __shfl_xor_sync(mask, sum, s);
  const CT inv = (CT)1 / sum;
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] *= inv;
  FrontVec<T, 4>::st(aw + tid * 4, v);
  const T* rf = ref + ((g / d.M) * d.L + l) * d.ref_dim;
  const CT r0 = to_acc<T>(rf[0]);
  const CT r1 = d.ref_dim == 2 ? to_acc<T>(rf[1]) : (CT)0;
  const CT tl = (CT)t_lvl[l];
  CT x0[4], x1[4];
  x0[0] = front_x<CT>(o[0], r0, r1, tl, (CT)d.P, d.ref_dim); x0[1] = (CT)0.5;
  x0[2] = front_x<CT>(o[1], r0, r1, tl, (CT)d.P, d.ref_dim); x0[3] = (CT)0.5;
  x1[0] = front_x<CT>(o[2], r0, r1, tl, (CT)d.P, d.ref_dim); x1[1] = (CT)0.5;
  x1[2] = front_x<CT>(o[3], r0, r1, tl, (CT)d.P, d.ref_dim); x1[3] = (CT)0.5;
  FrontVec<T, 4>::st(loc + tid * 8, x0);
  FrontVec<T, 4>::st(loc + tid * 8 + 4, x1);
}

template <typename T, bool SHFL>
__global__ void __launch_bounds__(kFrontThreads)
msda_frontend_bwd_lanes_kernel(const T* __restrict__ off, const T* __restrict__ aw,
                               const T* __restrict__ ref, const int64_t* __restrict__ t_lvl,
                               const T* __restrict__ gloc, const T* __restrict__ gaw,
                               T* __restrict__ goff, T* __restrict__ glogits, T* __restrict__ gref,
                               FrontDims d, int tpg) {
  using CT = typename AccOf<T>::type;
  const long long tid = (long long)blockIdx.x * kFrontThreads + threadIdx.x;
  const bool valid = tid < d.groups * tpg;
  const unsigned mask = __ballot_sync(0xffffffffu, valid);
  if (!valid) return;
  const long long g = tid / tpg;
  const int c = (int)(tid - g * tpg);
  const int ppl = d.P / 4;                        // pieces per level
  const int l = c / ppl;
  const long long nq = g / d.M;
  CT a[4], ga[4], o[4], t0[4], t1[4], gx[4];
  FrontVec<T, 4>::ld(aw + tid * 4, a);
  FrontVec<T, 4>::ld(gaw + tid * 4, ga);
  FrontVec<T, 4>::ld(off + tid * 4, o);
  FrontVec<T, 4>::ld(gloc + tid * 8, t0);
  FrontVec<T, 4>::ld(gloc + tid * 8 + 4, t1);
  gx[0] = t0[0]; gx[1] = t0[2]; gx[2] = t1[0]; gx[3] = t1[2];
  CT dot = (a[0] * ga[0] + a[1] * ga[1]) + (a[2] * ga[2] + a[3] * ga[3]);
  for (int s = tpg >> 1; s > 0; s >>= 1) dot += __shfl_xor_sync(mask, dot, s);
  const T* rf = ref + (nq * d.L + l) * d.ref_dim;
  const CT r1 = d.ref_dim == 2 ? to_acc<T>(rf[1]) : (CT)0;
  const CT tl = (CT)t_lvl[l];
  CT s0 = 0, s1 = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    a[k] = a[k] * (ga[k] - dot);
    s0 += gx[k];
    s1 += gx[k] * (CT)0.5 * (o[k] / (CT)d.P);
    gx[k] = d.ref_dim == 1 ? gx[k] / tl : gx[k] * (CT)0.5 * r1 / (CT)d.P;
  }
  FrontVec<T, 4>::st(glogits + tid * 4, a);
  FrontVec<T, 4>::st(goff + tid * 4, gx);
  if (gref != nullptr) {
    // the pieces of one level, then (SHFL) the heads of the query
    for (int s = ppl >> 1; s > 0; s >>= 1) {
      s0 += __shfl_xor_sync(mask, s0, s);
      s1 += __shfl_xor_sync(mask, s1, s);
    }
    T* gr = gref + (nq * d.L + l) * d.ref_dim;
    const bool first_piece = c - l * ppl == 0;
    if constexpr (SHFL) {
      for (int s = tpg; s < tpg * d.M; s <<= 1) {
        s0 += __shfl_xor_sync(mask, s0, s);
        s1 += __shfl_xor_sync(mask, s1, s);
      }
      if (first_piece && g - nq * d.M == 0) {
        gr[0] = from_acc<T, CT>(s0);
        if (d.ref_dim == 2) gr[1] = from_acc<T, CT>(s1);
      }
    } else if (first_piece) {
      front_atomic_add<T>(gr, s0);
      if (d.ref_dim == 2) front_atomic_add<T>(gr + 1, s1);
    }
  }
}


// ---- encoder reference points -------------------------------------------------
// DeformableTransformerEncoder.get_reference_points
// (models/deformable/unimodal_deformable_transformer.py:259-276; same code in the
// multimodal and sparse encoders): token t of level k, for every level l,
//   ref[n, s, l] = (t + 0.5) / (valid_ratio[n, k] * T_k) * valid_ratio[n, l]
// The reference builds it with a Python loop over a device tensor of level sizes
// (one host sync and four small kernels per level, every forward); here one
// thread per (batch item, token).
__global__ void __launch_bounds__(256)
msda_reference_points_kernel(const int64_t* __restrict__ t_lvl, const float* __restrict__ vr,
                             float* __restrict__ out, long long total, int S, int L) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const long long n = i / S;
  int t = (int)(i - n * S), k = 0, T = (int)t_lvl[0];
  while (k + 1 < L && t >= T) {
    t -= T;
    ++k;
    T = (int)t_lvl[k];
  }
  const float* v = vr + n * L;
  const float base = ((float)t + 0.5f) / (v[k] * (float)T);
  for (int l = 0; l < L; ++l) out[i * L + l] = base * v[l];
}

}  // namespace msda
