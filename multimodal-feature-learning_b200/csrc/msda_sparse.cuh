// msda_sparse.cuh -- token gather / scatter around the op in the sparse encoders
// (SURVEY.md section 8(f) row 3).
//
// models/sparse/unimodal_sparse_deformable_transformer.py:425-431 gathers the top-rho tokens
// (reference points, features, positions) with torch.gather over an index repeated C times, and
// :443-448 writes the layer's output back with one `output[i].scatter(...)` per batch item in a
// Python loop (sparse_token_nums differs per item).  Here: rows move as 16-byte vectors, one
// launch for the whole batch, the per-item token count read on the device.
//   gather : out[b, k, :] = x[b, idx[b, k], :]                       (k >= nums[b]: zeros)
//   scatter: out[b, s, :] = src[b, k, :] if idx[b, k] == s for a k < nums[b] else base[b, s, :]
//            through an inverse map inv[b, s] (-1 = keep base) built by msda_token_inverse_kernel,
//            so every output row is written exactly once and no row of `base` that is replaced
//            is read.
// Both are pure streaming kernels (HBM-bound): bytes = rows moved * C * e, read + written.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace msda {

__global__ void __launch_bounds__(256)
msda_token_inverse_kernel(const int64_t* __restrict__ idx, const int64_t* __restrict__ nums,
                          int* __restrict__ inv, int B, int S, int K) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long long)B * K) return;
  const int b = (int)(i / K), k = (int)(i - (long long)b * K);
  if (nums && k >= nums[b]) return;
  const int64_t s = idx[i];
  if (s >= 0 && s < S) inv[(long long)b * S + s] = k;
}

// rows of `row_bytes` bytes (a multiple of VEC); one warp per row, VEC bytes per lane and step
template <int VEC>
__device__ __forceinline__ void token_copy_row(char* __restrict__ dst, const char* __restrict__ src,
                                               int row_bytes, int lane) {
  for (int o = lane * VEC; o < row_bytes; o += 32 * VEC) {
    if (VEC == 16) {
      const uint4 v = src ? __ldcs(reinterpret_cast<const uint4*>(src + o)) : make_uint4(0, 0, 0, 0);
      __stcs(reinterpret_cast<uint4*>(dst + o), v);
    } else {
      dst[o] = src ? src[o] : (char)0;
    }
  }
}

template <int VEC>
__global__ void __launch_bounds__(256)
msda_token_gather_kernel(const char* __restrict__ x, const int64_t* __restrict__ idx,
                         const int64_t* __restrict__ nums, char* __restrict__ out, int B, int S, int K,
                         int row_bytes) {
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= (long long)B * K) return;
  const int b = (int)(row / K), k = (int)(row - (long long)b * K);
  const int64_t s = idx[row];
  const bool live = (!nums || k < nums[b]) && s >= 0 && s < S;
  token_copy_row<VEC>(out + row * row_bytes, live ? x + ((long long)b * S + s) * row_bytes : nullptr,
                      row_bytes, threadIdx.x & 31);
}

// ZERO_HIT: rows that the scatter replaced come out as zeros (the gradient w.r.t. `base`)
template <int VEC, bool ZERO_HIT>
__global__ void __launch_bounds__(256)
msda_token_scatter_kernel(const char* __restrict__ base, const int* __restrict__ inv,
                          const char* __restrict__ src, char* __restrict__ out, int B, int S, int K,
                          int row_bytes) {
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= (long long)B * S) return;
  const int b = (int)(row / S);
  const int k = inv[row];
  const char* from = k < 0 ? base + row * row_bytes
                           : (ZERO_HIT ? nullptr : src + ((long long)b * K + k) * row_bytes);
  token_copy_row<VEC>(out + row * row_bytes, from, row_bytes, threadIdx.x & 31);
}

}  // namespace msda
