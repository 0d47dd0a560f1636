// msda_dam.cuh -- decoder attention map: the attention weights of every sampling
// point scattered onto the flat token grid (utils/dam.py:20-73,
// attn_map_to_flat_grid; consumer models/criterion.py:246-283).  The same
// (location, weight) stream the op consumes, reduced over queries, levels' points:
//   t   = loc_x * T_l                       (no -0.5 here: dam.py:45)
//   out[bl, m, start_l + floor(t)]     += aw * (t - floor(t) - 1)      (dam.py:50, :57)
//   out[bl, m, start_l + floor(t) + 1] += aw * (t - floor(t))          (dam.py:49, :57)
// for indices inside [0, T_l).  (The factor of the low tap is negative, as in the
// reference: it keeps the 1-D remainder of a product of two negative margins.)
// One thread per sampling point, fp32 reds into the zero-filled [BL, M, S] grid
// (a few MB: L2-resident).
#pragma once
#include "msda_common.cuh"

namespace msda {

struct DamDims {
  long long points;   // BL * Q * M * L * P
  int Q, M, L, P, S;
  int loc_cols;       // 1: [.., P, 1] as the modules return it; 2: the op's (x, y) pairs
};

template <typename T>
__global__ void __launch_bounds__(256)
msda_dam_scatter_kernel(const T* __restrict__ loc, const T* __restrict__ aw,
                        const int64_t* __restrict__ t_lvl, const int64_t* __restrict__ start_lvl,
                        float* __restrict__ out, DamDims d) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= d.points) return;
  const long long lp = i / d.P;
  const int l = (int)(lp % d.L);
  const long long qm = lp / d.L;
  const int m = (int)(qm % d.M);
  const long long bl = qm / d.M / d.Q;
  const float x = (float)to_acc<T>(loc[i * d.loc_cols]);
  const float a = (float)to_acc<T>(aw[i]);
  const int T_l = (int)t_lvl[l];
  const float t = x * (float)T_l;
  const float t0 = floorf(t);
  if (!(t0 >= -1.f && t0 < (float)T_l)) return;   // both taps outside (or NaN)
  const int i0 = (int)t0;
  const float frac = t - t0;
  float* row = out + (bl * d.M + m) * d.S + (int)start_lvl[l];
  if (i0 >= 0) atomicAdd(row + i0, a * (frac - 1.f));
  if (i0 + 1 < T_l) atomicAdd(row + i0 + 1, a * frac);
}

}  // namespace msda
