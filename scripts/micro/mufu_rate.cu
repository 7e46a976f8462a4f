// Micro-benchmark (developer tool): MUFU.EX2 rate per SM sub-partition on this GPU, alone and fed by an FFMA, with 1..4
// warps per sub-partition.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, long long* clk, float c, float mc, int iters) {
  float v[64];
#pragma unroll
  for (int i = 0; i < 64; ++i) v[i] = (float)(threadIdx.x + i) * 1e-3f;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 64; ++i) {
      float x = v[i];
      if (MODE >= 1) x = fmaf(x, c, -mc);
      asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(x) : "f"(x));
      if (MODE == 2) x = x + v[(i + 1) & 63] * 1e-9f;
      v[i] = x;
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 64; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}

int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&clk, 8);
  const int iters = 200;
  for (int mode = 0; mode < 3; ++mode)
    for (int warps = 4; warps <= 16; warps *= 2) {
      long long h = 0;
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, warps * 32>>>(out, clk, 0.999f, 1e-4f, iters);
        if (mode == 1) k<1><<<148, warps * 32>>>(out, clk, 0.999f, 1e-4f, iters);
        if (mode == 2) k<2><<<148, warps * 32>>>(out, clk, 0.999f, 1e-4f, iters);
        cudaDeviceSynchronize();
      }
      cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost);
      const double per_sm_per_clk = (double)warps * 32 * 64 * iters / (double)h;
      printf("mode %d (0 ex2 only, 1 ffma+ex2, 2 ffma+ex2+ffma)  warps/SM %2d (%d per sub-partition): %lld clk, %.2f ex2/clk/SM, %.2f clk per warp-ex2 per sub-partition\n",
             mode, warps, warps / 4, h, per_sm_per_clk, 128.0 / per_sm_per_clk);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
