// Micro-benchmark (developer tool): issue rate of tcgen05.mma (M = 128, K = 16, fp16, both operands K-major
// SWIZZLE_128B in shared memory) when the A operand's sixteen 8-row groups are `pitch` rows apart and the operand starts
// `shift` rows into the tile -- the addressing conv_gemm.cu's haloed patches use (pitch 10, shift ky * 10 + kx) --
// against the canonical pitch 8 / shift 0.  One CTA, 4096 MMAs back to back, clock64 around them.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../../lipreading_video_generation_b200/csrc -o umma_rate_k umma_rate_k.cu -lcuda
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "umma.cuh"

using namespace b2d::umma;

__global__ void __launch_bounds__(128) k_rate(long long* out, int pitch, int shift, int N, int iters) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* Bm = smem;                         // [128 n][128 B]
  uint8_t* X = smem + 16384;                  // 256 rows x 128 B
  uint64_t* bar = (uint64_t*)(smem + 16384 + 256 * 128);
  uint32_t* slot = (uint32_t*)(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (16384 + 256 * 128) / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u;   // fp16 ones
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 128);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  const uint32_t idesc = make_idesc(128, N, 0, 0, 0, 0);
  if (tid == 0) {
    const uint32_t xa = smem_u32(X) + (uint32_t)shift * 128, ba = smem_u32(Bm);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        mma_f16_ss(tmem, make_sdesc_sw128(xa + k * 32, 16, (uint32_t)pitch * 128), make_sdesc_sw128(ba + k * 32, 16, 1024),
                   idesc, 1u);
    }
    mma_commit(bar);
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  const int smem = 16384 + 256 * 128 + 64 + 1024;
  cudaFuncSetAttribute(k_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 1024;
  const int cases[][2] = {{8, 0}, {8, 1}, {8, 4}, {10, 0}, {10, 1}, {10, 11}, {10, 22}, {12, 0}, {16, 0}, {16, 1}, {9, 0}};
  for (int N = 64; N <= 128; N += 64) {
    for (auto& c : cases) {
      k_rate<<<1, 128, smem>>>(d, c[0], c[1], N, 8);          // warm-up
      k_rate<<<1, 128, smem>>>(d, c[0], c[1], N, iters);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
      long long h;
      cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      printf("N = %3d  pitch %2d rows  shift %2d rows: %6.1f clk per MMA\n", N, c[0], c[1], (double)h / (iters * 4));
    }
  }
  return 0;
}
