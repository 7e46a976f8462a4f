// Microbenchmark: TMEM read bandwidth (tcgen05.ld 32x32b.x32) per SM with 4 and 8 reader warps.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_bw tmem_ld_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../lipreading_video_generation_b200/csrc/umma.cuh"
using namespace b2d::umma;

__global__ void k(int iters, long long* out, unsigned* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(&slot, 512);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  unsigned acc = 0;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    uint32_t r[128];
    tmem_ld32(base + 0, r); tmem_ld32(base + 32, r + 32); tmem_ld32(base + 64, r + 64); tmem_ld32(base + 96, r + 96);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 128; j += 16) acc += r[j];
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  if (acc == 0x12345678u) sink[0] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(slot, 512);
}

int main() {
  long long* d; unsigned* s; cudaMalloc(&d, 8 * 148); cudaMalloc(&s, 4);
  for (int warps : {4, 8}) {
    const int iters = 2000;
    k<<<148, warps * 32>>>(iters, d, s);
    cudaDeviceSynchronize();
    k<<<148, warps * 32>>>(iters, d, s);
    cudaError_t e = cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    double bytes = (double)warps * 32 * 128 * 4 * iters;
    printf("warps=%d err=%d cycles=%lld  bytes/clk/SM=%.1f  clk per 128x128 fp32 tile (4 warps' worth)=%.0f\n", warps, (int)e, h,
           bytes / h, (double)h / iters / (warps / 4));
  }
  return 0;
}
