// Micro-test (developer tool): may the start address of an MN-major SWIZZLE_128B shared-memory operand of tcgen05.mma
// be shifted by whole K rows (128 B each) that are NOT a multiple of the 8-row swizzle atom?  That is what lets the
// nine taps of a 3x3 weight-gradient be shifted views of ONE haloed X tile in shared memory.
//   D[m][n] = sum_{k<16} A[k][m] * X[s + k][n],  s = 0..15, with the descriptor's base-offset field 0 or (addr >> 7) & 7.
// The X tile is written the way TMA writes it (16-byte chunk index XOR (128-byte line index & 7), absolute in a
// 1024-byte aligned buffer).  Prints the number of mismatching accumulator elements per (shift, mode).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../../lipreading_video_generation_b200/csrc -o umma_shift umma_shift.cu -lcuda
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "umma.cuh"

using namespace b2d::umma;

__device__ __forceinline__ float a_val(int k, int m) { return (float)(((k * 5 + m * 3) % 7) - 3); }
__device__ __forceinline__ float x_val(int p, int n) { return (float)(((p * 7 + n * 11) % 13) - 6); }

__global__ void __launch_bounds__(128) k_shift(int* mismatches, int nshift) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* A = smem;            // two [16 rows][128 B] blocks (M = 0..63, 64..127), 2048 B apart
  uint8_t* X = smem + 8192;     // [64 rows][128 B]
  uint64_t* bar = (uint64_t*)(smem + 8192 + 64 * 128);
  uint32_t* slot = (uint32_t*)(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 16 * 128; i += 128) {
    const int k = i / 128, m = i % 128;
    const int blk = m >> 6, c = m & 63;
    const uint32_t off = (uint32_t)blk * 2048 + (uint32_t)k * 128 + (uint32_t)(((c >> 3) ^ (k & 7)) << 4) + (uint32_t)(c & 7) * 2;
    *(__half*)(A + off) = __float2half(a_val(k, m));
  }
  for (int i = tid; i < 64 * 64; i += 128) {
    const int p = i / 64, n = i % 64;
    const uint32_t off = (uint32_t)p * 128 + (uint32_t)(((n >> 3) ^ (p & 7)) << 4) + (uint32_t)(n & 7) * 2;
    *(__half*)(X + off) = __float2half(x_val(p, n));
  }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 64);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  const uint32_t idesc = make_idesc(128, 64, 1, 1, 0, 0);  // formats: 0 = f16
  uint32_t phase = 0;
  for (int s = 0; s < nshift; ++s) {
    for (int mode = 0; mode < 2; ++mode) {
      if (tid == 0) {
        const uint32_t xa = smem_u32(X) + (uint32_t)s * 128;
        uint64_t bd = make_sdesc_sw128(xa, 8192, 1024);
        if (mode == 1) bd |= (uint64_t)((xa >> 7) & 7u) << 49;
        mma_f16_ss(tmem, make_sdesc_sw128(smem_u32(A), 2048, 1024), bd, idesc, 0u);
        mma_commit(bar);
      }
      mbar_wait(bar, phase);
      phase ^= 1u;
      tc_fence_after();
      int bad = 0;
      const int m = tid;
      for (int c0 = 0; c0 < 64; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
        tmem_ld_wait();
        for (int j = 0; j < 16; ++j) {
          float want = 0.f;
          for (int k = 0; k < 16; ++k) want += a_val(k, m) * x_val(s + k, c0 + j);
          if (__uint_as_float(r[j]) != want) ++bad;
        }
      }
      atomicAdd(&mismatches[s * 2 + mode], bad);
      tc_fence_before();
      __syncthreads();
      tc_fence_after();
    }
  }
  if (warp == 0) tmem_dealloc(tmem, 64);
}

// Second question: may the two 64-element MN groups of an M = 128 operand be two OVERLAPPING shifted views of the same
// tile (LBO = a few rows, not a multiple of 1024 B)?   D[g * 64 + c][n] = sum_k X[s + g * lbo_rows + k][c] * B[k][n]
__global__ void __launch_bounds__(128) k_pair(int* mismatches, int nshift, int lbo_rows) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* Bm = smem;           // [16 rows][128 B] (N = 0..63)
  uint8_t* X = smem + 8192;     // [128 rows][128 B]
  uint64_t* bar = (uint64_t*)(smem + 8192 + 128 * 128);
  uint32_t* slot = (uint32_t*)(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 16 * 64; i += 128) {
    const int k = i / 64, c = i % 64;
    const uint32_t off = (uint32_t)k * 128 + (uint32_t)(((c >> 3) ^ (k & 7)) << 4) + (uint32_t)(c & 7) * 2;
    *(__half*)(Bm + off) = __float2half(a_val(k, c));
  }
  for (int i = tid; i < 128 * 64; i += 128) {
    const int p = i / 64, n = i % 64;
    const uint32_t off = (uint32_t)p * 128 + (uint32_t)(((n >> 3) ^ (p & 7)) << 4) + (uint32_t)(n & 7) * 2;
    *(__half*)(X + off) = __float2half(x_val(p, n));
  }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 64);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  const uint32_t idesc = make_idesc(128, 64, 1, 1, 0, 0);
  uint32_t phase = 0;
  for (int s = 0; s < nshift; ++s) {
    if (tid == 0) {
      const uint32_t xa = smem_u32(X) + (uint32_t)s * 128;
      mma_f16_ss(tmem, make_sdesc_sw128(xa, (uint32_t)lbo_rows * 128, 1024), make_sdesc_sw128(smem_u32(Bm), 8192, 1024),
                 idesc, 0u);
      mma_commit(bar);
    }
    mbar_wait(bar, phase);
    phase ^= 1u;
    tc_fence_after();
    int bad = 0;
    const int g = tid >> 6, c = tid & 63;
    for (int c0 = 0; c0 < 64; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
      tmem_ld_wait();
      for (int j = 0; j < 16; ++j) {
        float want = 0.f;
        for (int k = 0; k < 16; ++k) want += x_val(s + g * lbo_rows + k, c) * a_val(k, c0 + j);
        if (__uint_as_float(r[j]) != want) ++bad;
      }
    }
    atomicAdd(&mismatches[s], bad);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  if (warp == 0) tmem_dealloc(tmem, 64);
}

int main() {
  const int nshift = 17;
  int* d;
  cudaMalloc(&d, nshift * 2 * sizeof(int));
  cudaMemset(d, 0, nshift * 2 * sizeof(int));
  const int smem = 8192 + 64 * 128 + 64 + 1024;
  cudaFuncSetAttribute(k_shift, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k_shift<<<1, 128, smem>>>(d, nshift);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  int h[nshift * 2];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("shift(rows)  mismatches[base_offset=0]  mismatches[base_offset=(addr>>7)&7]   (of 8192)\n");
  for (int s = 0; s < nshift; ++s) printf("%5d %12d %12d\n", s, h[s * 2], h[s * 2 + 1]);
  const int smem2 = 8192 + 128 * 128 + 64 + 1024;
  cudaFuncSetAttribute(k_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2);
  const int lbos[4] = {0, 1, 64, 66};
  for (int li = 0; li < 4; ++li) {
    cudaMemset(d, 0, nshift * 2 * sizeof(int));
    k_pair<<<1, 128, smem2>>>(d, nshift, lbos[li]);
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("pair views, LBO = %d rows: mismatches per shift:", lbos[li]);
    for (int s2 = 0; s2 < nshift; ++s2) printf(" %d", h[s2]);
    printf("\n");
  }
  return 0;
}
