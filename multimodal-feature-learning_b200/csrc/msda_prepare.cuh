// msda_prepare.cuh -- prepare_encoder_inputs in one pass (second half).
//
// models/deformable/unimodal_deformable_transformer.py:90-134 (and its copies in the multimodal and
// sparse transformers): per level transpose [B, C, T_l] -> [B, T_l, C], add the level embedding to
// the positional embedding, concatenate the levels, concatenate the masks, valid ratio per level --
// a Python loop of 2 transposes + 1 add per level, 3 cats and L small reductions.  Here: one
// tiled-transpose kernel writes src_flatten, lvl_pos_embed_flatten and mask_flatten directly at
// their level offsets (reads coalesced along t, writes coalesced along c, 32 x 33 shared tile), and
// one warp per (batch, level) counts the valid tokens.  Streaming, HBM-bound:
// bytes = 2 * B * S * C * e read + the same written.
#pragma once
#include "msda_common.cuh"

namespace msda {

constexpr int kPrepMaxL = 8;

struct PrepLevels {
  const void* src[kPrepMaxL];
  const void* pos[kPrepMaxL];
  const unsigned char* mask[kPrepMaxL];
  int T[kPrepMaxL];
  int start[kPrepMaxL];
  int tile0[kPrepMaxL + 1];   // first t-tile of the level in the grid's x dimension
  int L;
};

template <typename T>
__global__ void __launch_bounds__(256)
msda_prepare_encoder_kernel(PrepLevels lv, const T* __restrict__ level_embed, T* __restrict__ out_src,
                            T* __restrict__ out_pos, unsigned char* __restrict__ out_mask, int B, int C,
                            int S) {
  __shared__ T ts[32][33], tp[32][33];
  int l = 0;
  while (l + 1 < lv.L && (int)blockIdx.x >= lv.tile0[l + 1]) ++l;
  const int t0 = ((int)blockIdx.x - lv.tile0[l]) * 32, c0 = blockIdx.y * 32, b = blockIdx.z;
  const int Tl = lv.T[l], tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const T* s = static_cast<const T*>(lv.src[l]) + (size_t)b * C * Tl;
  const T* p = static_cast<const T*>(lv.pos[l]) + (size_t)b * C * Tl;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = c0 + ty + 8 * k, t = t0 + tx;
    if (c < C && t < Tl) {
      ts[ty + 8 * k][tx] = s[(size_t)c * Tl + t];
      tp[ty + 8 * k][tx] = p[(size_t)c * Tl + t];
    }
  }
  __syncthreads();
  const int c = c0 + tx;
  using CT = typename AccOf<T>::type;
  const CT le = c < C ? to_acc<T>(level_embed[(size_t)l * C + c]) : (CT)0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int t = t0 + ty + 8 * k;
    if (c < C && t < Tl) {
      const size_t o = ((size_t)b * S + lv.start[l] + t) * C + c;
      out_src[o] = ts[tx][ty + 8 * k];
      out_pos[o] = from_acc<T, CT>(to_acc<T>(tp[tx][ty + 8 * k]) + le);
    }
  }
  if (blockIdx.y == 0 && threadIdx.x < 32 && t0 + tx < Tl)
    out_mask[(size_t)b * S + lv.start[l] + t0 + tx] = lv.mask[l][(size_t)b * Tl + t0 + tx];
}

// valid_ratios[b, l] = (number of tokens of level l of item b that are not padding) / T_l
__global__ void __launch_bounds__(32)
msda_valid_ratio_kernel(PrepLevels lv, float* __restrict__ out, int B) {
  const int l = blockIdx.x, b = blockIdx.y, Tl = lv.T[l];
  int n = 0;
  for (int t = threadIdx.x; t < Tl; t += 32) n += lv.mask[l][(size_t)b * Tl + t] ? 0 : 1;
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) n += __shfl_xor_sync(0xffffffffu, n, s);
  // count * (1 / T_l): torch divides a tensor by a Python scalar this way (get_valid_ratio, :86)
  if (threadIdx.x == 0) out[(size_t)b * lv.L + l] = (float)n * (1.0f / (float)Tl);
}

}  // namespace msda
