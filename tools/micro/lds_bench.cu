// Microbenchmark: shared-memory wavefront cost of the access patterns the tile kernels use.
// nvcc -arch=sm_100a -O3 -o lds_bench lds_bench.cu && ./lds_bench
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
  __shared__ __align__(16) float s[8192];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) s[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int off;   // in floats
  if (MODE == 0) off = lane * 4;                 // LDS.128 all distinct (512 B)
  else if (MODE == 1) off = (lane >> 3) * 192;   // LDS.128 4 groups, same bank, 768 B apart (our Wt)
  else if (MODE == 2) off = 0;                   // LDS.128 all lanes same address
  else if (MODE == 3) off = (lane >> 3) * 4;     // LDS.128 4 groups, contiguous 64 B
  else if (MODE == 4) off = (lane >> 3) * 196;   // LDS.128 4 groups, skewed banks
  else if (MODE == 5) off = (lane & 7) * 4;      // LDS.128 8 distinct x 16 B (128 B), 4 groups same (warp-wide V row)
  else if (MODE == 6) off = lane * 2;            // LDS.64 distinct (256 B)
  else if (MODE == 7) off = (lane >> 3);         // LDS.32 4 distinct words
  else off = (lane & 7) * 4 + (lane >> 3) * 64;  // LDS.128 4 groups x 128 B distinct rows (block-private V rows)
  float4 acc = make_float4(0, 0, 0, 0);
  const float* p = s + off;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      const int o = ((i + u) & 7) * 256;
      if (MODE == 6) { float2 v = *reinterpret_cast<const float2*>(p + o); acc.x += v.x; acc.y += v.y; }
      else if (MODE == 7) { float v = p[o]; acc.x += v; }
      else { float4 v = *reinterpret_cast<const float4*>(p + o); acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
}

template <int MODE> void run(const char* name, int warps) {
  float* out; long long* cyc;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 1024);
  const int iters = 2000;
  k<MODE><<<1, warps * 32>>>(out, iters, cyc);
  k<MODE><<<1, warps * 32>>>(out, iters, cyc);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-55s warps %2d: %.2f cycles per warp-LDS (SM-wide %.2f per LDS)\n", name, warps, (double)h / (iters * 16), (double)h / (iters * 16) / warps);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int w : {8, 16}) {
    if (w == 8) {
      run<0>("LDS.128 32 distinct (512 B)", 8); run<1>("LDS.128 4 groups same bank (Wt now)", 8);
      run<2>("LDS.128 uniform address", 8); run<3>("LDS.128 4 groups contiguous 64 B", 8);
      run<4>("LDS.128 4 groups skewed", 8); run<5>("LDS.128 one 128 B row for all 4 groups", 8);
      run<6>("LDS.64 32 distinct (256 B)", 8); run<7>("LDS.32 4 distinct words", 8);
      run<8>("LDS.128 4 groups x 128 B rows", 8);
    } else {
      run<0>("LDS.128 32 distinct (512 B)", 16); run<1>("LDS.128 4 groups same bank (Wt now)", 16);
      run<2>("LDS.128 uniform address", 16); run<3>("LDS.128 4 groups contiguous 64 B", 16);
      run<4>("LDS.128 4 groups skewed", 16); run<5>("LDS.128 one 128 B row for all 4 groups", 16);
      run<6>("LDS.64 32 distinct (256 B)", 16); run<7>("LDS.32 4 distinct words", 16);
      run<8>("LDS.128 4 groups x 128 B rows", 16);
    }
  }
  return 0;
}
