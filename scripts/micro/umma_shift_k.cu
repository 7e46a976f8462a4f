// Micro-test (developer tool), companion of umma_shift.cu for K-MAJOR operands: may the M = 128 rows of a K-major
// SWIZZLE_128B shared-memory operand of tcgen05.mma be 16 groups of 8 rows that are `pitch` rows apart (stride byte
// offset = pitch * 128 B, not a multiple of 1024), starting at ANY 128-byte row of a tile that TMA wrote into a
// 1024-byte aligned buffer?  That is what lets the nine taps of a 3x3 convolution's forward / data-gradient pass be
// shifted views of ONE haloed activation tile: a tile of 16 image rows x 8 pixels, loaded with its halo as
// (16 + 2) x (8 + 2) pixels x 64 channels, has the 8 pixels of image row g at tile row (g + ky) * 10 + kx.
//   D[m][n] = sum_{k<16} X[s + (m / 8) * pitch + m % 8][kk * 16 + k] * B[n][kk * 16 + k]
// for s = 0..2*pitch+2, kk = 0..3, pitch in {8, 10, 12}.  Prints the number of mismatching accumulator elements.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I../../lipreading_video_generation_b200/csrc -o umma_shift_k umma_shift_k.cu -lcuda
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "umma.cuh"

using namespace b2d::umma;

__device__ __forceinline__ float b_val(int n, int k) { return (float)(((k * 5 + n * 3) % 7) - 3); }
__device__ __forceinline__ float x_val(int p, int c) { return (float)(((p * 7 + c * 11) % 13) - 6); }

constexpr int XROWS = 232;      // 23 x 1024 B would hold 184 rows; room for pitch 12: 26 + 15 * 12 + 7 = 213

__global__ void __launch_bounds__(128) k_patch(int* mismatches, int pitch, int nshift) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* Bm = smem;                         // [64 n][128 B], K-major
  uint8_t* X = smem + 8192;                   // [XROWS pixels][128 B = 64 channels], K-major
  uint64_t* bar = (uint64_t*)(smem + 8192 + XROWS * 128);
  uint32_t* slot = (uint32_t*)(bar + 1);
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 64 * 64; i += 128) {
    const int n = i / 64, k = i % 64;
    const uint32_t off = (uint32_t)n * 128 + (uint32_t)(((k >> 3) ^ (n & 7)) << 4) + (uint32_t)(k & 7) * 2;
    *(__half*)(Bm + off) = __float2half(b_val(n, k));
  }
  for (int i = tid; i < XROWS * 64; i += 128) {
    const int p = i / 64, c = i % 64;
    const uint32_t off = (uint32_t)p * 128 + (uint32_t)(((c >> 3) ^ (p & 7)) << 4) + (uint32_t)(c & 7) * 2;
    *(__half*)(X + off) = __float2half(x_val(p, c));
  }
  if (tid == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 64);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  const uint32_t idesc = make_idesc(128, 64, 0, 0, 0, 0);  // both operands K-major, f16
  uint32_t phase = 0;
  for (int s = 0; s < nshift; ++s) {
    for (int kk = 0; kk < 4; ++kk) {
      if (tid == 0) {
        const uint32_t xa = smem_u32(X) + (uint32_t)s * 128 + (uint32_t)kk * 32;
        mma_f16_ss(tmem, make_sdesc_sw128(xa, 16, (uint32_t)pitch * 128), make_sdesc_sw128(smem_u32(Bm) + kk * 32, 16, 1024),
                   idesc, 0u);
        mma_commit(bar);
      }
      mbar_wait(bar, phase);
      phase ^= 1u;
      tc_fence_after();
      int bad = 0;
      const int m = tid;
      const int xr = s + (m >> 3) * pitch + (m & 7);
      for (int c0 = 0; c0 < 64; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
        tmem_ld_wait();
        for (int j = 0; j < 16; ++j) {
          float want = 0.f;
          for (int k = 0; k < 16; ++k) want += x_val(xr, kk * 16 + k) * b_val(c0 + j, kk * 16 + k);
          if (__uint_as_float(r[j]) != want) ++bad;
        }
      }
      atomicAdd(&mismatches[s * 4 + kk], bad);
      tc_fence_before();
      __syncthreads();
      tc_fence_after();
    }
  }
  if (warp == 0) tmem_dealloc(tmem, 64);
}

int main() {
  const int maxshift = 32;
  int* d;
  cudaMalloc(&d, maxshift * 4 * sizeof(int));
  const int smem = 8192 + XROWS * 128 + 64 + 1024;
  cudaFuncSetAttribute(k_patch, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int pitches[3] = {8, 10, 12};
  for (int pi = 0; pi < 3; ++pi) {
    const int pitch = pitches[pi], nshift = 2 * pitch + 3;
    cudaMemset(d, 0, maxshift * 4 * sizeof(int));
    k_patch<<<1, 128, smem>>>(d, pitch, nshift);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    int h[maxshift * 4];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    int total = 0;
    printf("pitch %2d rows (SBO = %4d B): mismatches (of 8192) per shift, summed over the four K = 16 slices:", pitch, pitch * 128);
    for (int s = 0; s < nshift; ++s) { const int b = h[s * 4] + h[s * 4 + 1] + h[s * 4 + 2] + h[s * 4 + 3]; total += b; printf(" %d", b); }
    printf("  | total %d\n", total);
  }
  return 0;
}
