// common.cuh -- shared host/device helpers for libb200diff (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/b2d.h"

namespace b2d {

// thread-local error text returned by b2d_last_error()
char* err_buf();
int set_err(int code, const char* fmt, ...);

#define B2D_CHECK_CUDA(expr)                                                   \
  do {                                                                         \
    cudaError_t _e = (expr);                                                   \
    if (_e != cudaSuccess)                                                     \
      return b2d::set_err(B2D_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,        \
                          cudaGetErrorString(_e), __FILE__, __LINE__);         \
  } while (0)

#define B2D_CHECK_LAUNCH()                                                     \
  do {                                                                         \
    b2d::count_launch();                                                       \
    cudaError_t _e = cudaGetLastError();                                       \
    if (_e != cudaSuccess)                                                     \
      return b2d::set_err(B2D_ERR_CUDA, "kernel launch failed: %s (%s:%d)",    \
                          cudaGetErrorString(_e), __FILE__, __LINE__);         \
  } while (0)

#define B2D_REQUIRE(cond, code, ...)                                           \
  do {                                                                         \
    if (!(cond)) return b2d::set_err(code, __VA_ARGS__);                       \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int num_sms();  // cached SM count of the current device (148 on B200)
void count_launch();  // bumps the counter returned by b2d_launch_count()

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream_f4(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ldg_nc_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&v);
  return __bfloat1622float2(h);
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }

}  // namespace b2d
