// msda_det.cuh -- the deterministic backward (MSDA_FLAG_DETERMINISTIC).
//
// grad_sampling_loc and grad_attn_weight are reproducible in every kernel family
// (each point is reduced by its owner in a fixed order); grad_value is not: its
// contributions meet in floating-point atomics whose order changes from run to
// run (as in the reference, .cuh:407-511).  Instead of sorting the 2 * 10^7
// (row, contribution) pairs of an encoder call into a fixed order, this variant
// makes the order irrelevant: every contribution is rounded ONCE to a 64-bit
// fixed-point number and accumulated with integer atomics, which commute, so the
// sum is bitwise the same whatever the schedule.  The scale 2^e is chosen on the
// device from max|grad_out| * max|attn_weight| (a bound on any contribution; tap
// weights are <= 1) and the largest possible number of addends per element,
// Lq * L * P * 4, so that no sum can reach 2^62; the quantum, bound * 2^-(61 - log2
// addends), is ~2^-20 of an fp32 ulp of the bound.
//
// Workspace: a 32-byte header, then one int64 per grad_value element.
#pragma once
#include "msda_common.cuh"

namespace msda {

struct DetHeader {
  unsigned amax_go, amax_aw;   // bit patterns of max|x| as fp32 (>= 0: ordered like integers)
  int e;                       // scale = 2^e
  int nonfinite;               // max|grad_out| * max|attn_weight| is NaN or inf: grad_value becomes NaN
  double scale, inv_scale;
};
static_assert(sizeof(DetHeader) == 32, "workspace layout");

template <typename T> __device__ __forceinline__ float det_abs_up(T v) { return fabsf((float)to_acc<T>(v)); }
template <> __device__ __forceinline__ float det_abs_up<double>(double v) { return __double2float_ru(fabs(v)); }

template <typename T>
__global__ void __launch_bounds__(256)
msda_det_amax_kernel(const T* __restrict__ x, size_t n, unsigned* __restrict__ out) {
  float mx = 0.f;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
    const float a = det_abs_up<T>(x[i]);
    mx = a > mx || a != a ? a : mx;          // a NaN sticks: the header then says "not finite"
  }
  unsigned b = __float_as_uint(mx);
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) b = max(b, __shfl_xor_sync(0xffffffffu, b, s));
  if ((threadIdx.x & 31) == 0 && b != 0) atomicMax(out, b);   // max commutes: deterministic
}

__global__ void msda_det_scale_kernel(DetHeader* h, int log2_addends) {
  const float bound = __uint_as_float(h->amax_go) * __uint_as_float(h->amax_aw);
  int e = 0;
  if (bound > 0.f && bound <= 3.0e38f) {      // finite, non-zero (NaN fails both)
    e = 61 - (ilogbf(bound) + 1) - log2_addends;
    e = e > 120 ? 120 : e < -120 ? -120 : e;
  }
  h->e = e;
  h->nonfinite = !(bound <= 3.0e38f);     // NaN or inf (the reference's float atomics propagate them too)
  h->scale = ldexp(1.0, e);
  h->inv_scale = ldexp(1.0, -e);
}

// grad_value accumulation policies of the generic backward kernel
template <typename GV, typename CT> struct AccAdd {
  static __device__ __forceinline__ void add(GV* p, CT v, CT) { atomicAdd(p, (GV)v); }
};
template <> struct AccAdd<long long, float> {
  static __device__ __forceinline__ void add(long long* p, float v, float scale) {
    atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)__float2ll_rn(v * scale));
  }
};
template <> struct AccAdd<long long, double> {
  static __device__ __forceinline__ void add(long long* p, double v, double scale) {
    atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)__double2ll_rn(v * scale));
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
msda_det_convert_kernel(const long long* __restrict__ acc, T* __restrict__ gv, size_t n,
                        const DetHeader* __restrict__ h) {
  using CT = typename AccOf<T>::type;
  const double inv = h->nonfinite ? __longlong_as_double(0x7ff8000000000000ll) : h->inv_scale;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
    gv[i] = from_acc<T, CT>((CT)((double)acc[i] * inv));
}

}  // namespace msda
