// common.cuh -- shared host/device helpers for libb200diff (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <mutex>
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

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) exactly once per (kernel, device), race-free (lazily
// initialised immutable kernel attributes guarded by std::call_once).  Use inside an int-returning entry point.
#define B2D_SET_SMEM_ONCE(kernel, bytes)                                                              \
  do {                                                                                                \
    static std::once_flag _once[16];                                                                  \
    static cudaError_t _err[16];                                                                      \
    int _dev = 0;                                                                                     \
    cudaGetDevice(&_dev);                                                                             \
    _dev &= 15;                                                                                       \
    std::call_once(_once[_dev], [&] {                                                                 \
      _err[_dev] = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)); \
    });                                                                                               \
    B2D_CHECK_CUDA(_err[_dev]);                                                                       \
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
// fire-and-forget fp32 x4 accumulation into global memory (REDG.E.ADD.F32x4): one L2 operation for 16 bytes
__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
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
// Storage type: every 16-bit tensor is fp16 (11-bit significand) with fp32 accumulation.
// Forward values are O(1) after GroupNorm so the range is safe and conversions saturate; gradient tensors
// carry a power-of-two loss scale chosen on the device per step (b2d_grad_scale) and the fp32 parameter
// gradients are unscaled at the end (b2d_scale_f32).  bf16 storage was measured to miss north_star's own
// 2e-2 gradient tolerance on this network (profiles/parity_r1_status.md); mixing f16 x bf16 operands in one
// tcgen05.mma kind::f16 raises an illegal-instruction fault on sm_100a (measured), so one format is used.
constexpr int FMT_F16 = 0, FMT_BF16 = 1;
__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {   // saturating: one F2FP.SATFINITE
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack_f16x2_nosat(float lo, float hi) {   // values known to be in [0, 1]
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_f16x2(uint32_t v) {
  __half2 h = *reinterpret_cast<__half2*>(&v);
  return __half22float2(h);
}
// fmt-dispatched pair conversion (fmt is warp-uniform)
__device__ __forceinline__ uint32_t pack_x2(float lo, float hi, int fmt) {
  return fmt == FMT_F16 ? pack_f16x2(lo, hi) : pack_bf16x2(lo, hi);
}
__device__ __forceinline__ float2 unpack_x2(uint32_t v, int fmt) {
  return fmt == FMT_F16 ? unpack_f16x2(v) : unpack_bf16x2(v);
}
// 8 consecutive 16-bit elements <-> 8 floats
__device__ __forceinline__ void load8_fmt(const void* p, float* v, int fmt) {
  uint4 u = ldg_nc_u4(p);
  float2 a = unpack_x2(u.x, fmt), b = unpack_x2(u.y, fmt), c = unpack_x2(u.z, fmt), d = unpack_x2(u.w, fmt);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void store8_fmt(void* p, const float* v, int fmt) {
  uint4 u;
  u.x = pack_x2(v[0], v[1], fmt); u.y = pack_x2(v[2], v[3], fmt);
  u.z = pack_x2(v[4], v[5], fmt); u.w = pack_x2(v[6], v[7], fmt);
  *reinterpret_cast<uint4*>(p) = u;
}

// 8 packed fp16 -> fp32
__device__ __forceinline__ void unpack8(const uint4& u, float* v) {
  float2 a = unpack_f16x2(u.x), b = unpack_f16x2(u.y), c = unpack_f16x2(u.z), d = unpack_f16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}

// one MUFU.EX2 (exp2f() adds a denormal-range pre/post scale: FSETP + 2 FMUL per element)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// sigmoid with one MUFU.EX2 and one MUFU.RCP.  (1.0f / (1.0f + __expf(-z)) compiles to an IEEE division with a
// slow-path branch plus the denormal-range rescale of ex2: ~26 instructions per element measured in gn_apply,
// which made the GroupNorm kernels issue-bound at 65 % instead of HBM-bound; profiles/r1_groupnorm.md.)
__device__ __forceinline__ float fast_sigmoid(float z) {
  float r;
  const float e = fast_exp2(-1.4426950408889634f * z);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return r;
}

__device__ __forceinline__ float silu_f(float x) { return x * fast_sigmoid(x); }

// Philox-4x32-10 keyed by (seed, element index / 4): dropout masks reproducible in the backward
// (GroupNorm dropout, short-sequence attention dropout).
__device__ __forceinline__ uint4 philox4(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1) {
  uint32_t x0 = c0, x1 = c1, x2 = 0x243F6A88u, x3 = 0x85A308D3u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, x0), lo0 = 0xD2511F53u * x0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, x2), lo1 = 0xCD9E8D57u * x2;
    const uint32_t y0 = hi1 ^ x1 ^ k0, y2 = hi0 ^ x3 ^ k1;
    x0 = y0; x1 = lo1; x2 = y2; x3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(x0, x1, x2, x3);
}

}  // namespace b2d
