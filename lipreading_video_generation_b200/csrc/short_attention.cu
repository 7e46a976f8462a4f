// short_attention.cu -- self-attention over very short sequences (L <= 16 tokens, head dim 64): the twelve
// transformer layers of the audio encoder the reference instantiates in unet_audio.py:10-18 (HF transformers
// `Wav2Vec2Attention`, third-party) see 4000-sample windows = 12 frames, 12 heads.  The library flash kernels tile
// that 12 x 12 problem as 128 x 128 (0.067 ms forward + 0.134 ms backward per layer at 256 windows, plus the dropout
// mask kernels); here one warp owns one (window, head): Q, K, V rows sit in shared memory as fp32, the L x L score
// matrix too, softmax + dropout are done by one lane per row, and nothing but q, k, v (and d_out) is read from HBM.
// The backward recomputes the probabilities and the Philox keep-mask instead of saving them.
//
//   out[b, i, h, :] = sum_j drop(softmax_j(scale * q_i . k_j)) v_j       layouts: q/k/v (B, H, L, 64) views with
//   strides (sb, sh, sl, 1) of the Linear outputs; out / d_out / dq / dk / dv contiguous (B, L, H, 64).
#include "common.cuh"
#include <cuda_bf16.h>

namespace b2d {
namespace {

constexpr int SA_LMAX = 16;   // tokens
constexpr int SA_D = 64;      // head dim: a lane owns columns lane and lane + 32
constexpr int SA_RS = 65;     // fp32 row stride of the staged rows (odd: rows land in different banks)
constexpr int SA_LP = 17;     // row stride of the L x L matrices
constexpr int SA_WARPS = 4;   // (window, head) pairs per block

struct SaParams {
  const void *q, *k, *v;
  long long sb, sl, sh;       // element strides of q / k / v: batch, token, head
  int B, H, L;
  float scale, drop_p;
  unsigned long long seed;
};

template <typename T> __device__ __forceinline__ float ldf(const T* p) { return (float)*p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float ldf<__half>(const __half* p) { return __half2float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void stf<__half>(__half* p, float v) { *p = __float2half_rn(v); }

// keep-mask of probability row i of pair bh (the same call in forward and backward)
__device__ __forceinline__ void row_keep(unsigned long long seed, long long bh, int i, float p, bool* keep) {
  const uint32_t thr = (uint32_t)(p * 65536.0f);
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const unsigned long long ctr = ((unsigned long long)bh * SA_LMAX + i) * 2 + half;
    const uint4 a = philox4((uint32_t)seed, (uint32_t)(seed >> 32), (uint32_t)ctr, (uint32_t)(ctr >> 32));
    const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      keep[half * 8 + 2 * t] = (w[t] & 0xFFFFu) >= thr;
      keep[half * 8 + 2 * t + 1] = (w[t] >> 16) >= thr;
    }
  }
}

template <typename T>
__device__ __forceinline__ void stage_rows(const T* base, long long sl, int L, float* dst, int lane) {
  for (int i = 0; i < L; ++i) {
    dst[i * SA_RS + lane] = ldf(base + i * sl + lane);
    dst[i * SA_RS + lane + 32] = ldf(base + i * sl + lane + 32);
  }
}

// dst[i][j] = mul * a_i . b_j, one (i, j) pair per lane and round
__device__ __forceinline__ void dot_pairs(const float* a, const float* b, float* dst, int L, float mul, int lane) {
  for (int p = lane; p < L * L; p += 32) {
    const int i = p / L, j = p - i * L;
    const float* ra = a + i * SA_RS;
    const float* rb = b + j * SA_RS;
    float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
#pragma unroll
    for (int d = 0; d < SA_D; d += 4) {
      c0 = fmaf(ra[d], rb[d], c0);
      c1 = fmaf(ra[d + 1], rb[d + 1], c1);
      c2 = fmaf(ra[d + 2], rb[d + 2], c2);
      c3 = fmaf(ra[d + 3], rb[d + 3], c3);
    }
    dst[i * SA_LP + j] = ((c0 + c1) + (c2 + c3)) * mul;
  }
}

// in-place softmax of row `lane` of ss (lanes >= L idle)
__device__ __forceinline__ void softmax_row(float* ss, int L, int lane) {
  if (lane < L) {
    float* row = ss + lane * SA_LP;
    float m = row[0];
    for (int j = 1; j < L; ++j) m = fmaxf(m, row[j]);
    float sum = 0.f;
    for (int j = 0; j < L; ++j) {
      const float e = __expf(row[j] - m);
      row[j] = e;
      sum += e;
    }
    const float inv = 1.0f / sum;
    for (int j = 0; j < L; ++j) row[j] *= inv;
  }
}

template <typename T>
__global__ void __launch_bounds__(SA_WARPS * 32)
short_attention_fwd_kernel(SaParams p, T* __restrict__ out) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long bh = (long long)blockIdx.x * SA_WARPS + warp;
  if (bh >= (long long)p.B * p.H) return;
  const int b = (int)(bh / p.H), h = (int)(bh - (long long)b * p.H);
  const int L = p.L;
  float* sq = smem + warp * (2 * SA_LMAX * SA_RS + SA_LMAX * SA_LP);
  float* sk = sq + SA_LMAX * SA_RS;
  float* ss = sk + SA_LMAX * SA_RS;
  const long long off = b * p.sb + h * p.sh;
  stage_rows((const T*)p.q + off, p.sl, L, sq, lane);
  stage_rows((const T*)p.k + off, p.sl, L, sk, lane);
  float v0[SA_LMAX], v1[SA_LMAX];
#pragma unroll
  for (int j = 0; j < SA_LMAX; ++j) {
    v0[j] = j < L ? ldf((const T*)p.v + off + j * p.sl + lane) : 0.f;
    v1[j] = j < L ? ldf((const T*)p.v + off + j * p.sl + lane + 32) : 0.f;
  }
  __syncwarp();
  dot_pairs(sq, sk, ss, L, p.scale, lane);
  __syncwarp();
  softmax_row(ss, L, lane);
  if (p.drop_p > 0.f && lane < L) {
    bool keep[SA_LMAX];
    row_keep(p.seed, bh, lane, p.drop_p, keep);
    const float ks = 1.0f / (1.0f - p.drop_p);
#pragma unroll
    for (int j = 0; j < SA_LMAX; ++j)
      if (j < L) ss[lane * SA_LP + j] = keep[j] ? ss[lane * SA_LP + j] * ks : 0.f;
  }
  __syncwarp();
  for (int i = 0; i < L; ++i) {
    float o0 = 0.f, o1 = 0.f;
#pragma unroll
    for (int j = 0; j < SA_LMAX; ++j) {
      if (j < L) {
        const float pd = ss[i * SA_LP + j];
        o0 = fmaf(pd, v0[j], o0);
        o1 = fmaf(pd, v1[j], o1);
      }
    }
    T* o = out + (((long long)b * L + i) * p.H + h) * SA_D;
    stf(o + lane, o0);
    stf(o + lane + 32, o1);
  }
}

template <typename T>
__global__ void __launch_bounds__(SA_WARPS * 32)
short_attention_bwd_kernel(SaParams p, const T* __restrict__ dout, T* __restrict__ dq, T* __restrict__ dk,
                           T* __restrict__ dv) {
  extern __shared__ float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long bh = (long long)blockIdx.x * SA_WARPS + warp;
  if (bh >= (long long)p.B * p.H) return;
  const int b = (int)(bh / p.H), h = (int)(bh - (long long)b * p.H);
  const int L = p.L;
  float* sq = smem + warp * (4 * SA_LMAX * SA_RS + 2 * SA_LMAX * SA_LP);
  float* sk = sq + SA_LMAX * SA_RS;
  float* sv = sk + SA_LMAX * SA_RS;
  float* sdo = sv + SA_LMAX * SA_RS;
  float* ss = sdo + SA_LMAX * SA_RS;      // P, then dropped P
  float* sd = ss + SA_LMAX * SA_LP;       // d(dropped P), then scale * dS
  const long long off = b * p.sb + h * p.sh;
  const long long ooff = ((long long)b * L * p.H + h) * SA_D;     // (B, L, H, 64): token stride H * 64
  const long long osl = (long long)p.H * SA_D;
  stage_rows((const T*)p.q + off, p.sl, L, sq, lane);
  stage_rows((const T*)p.k + off, p.sl, L, sk, lane);
  stage_rows((const T*)p.v + off, p.sl, L, sv, lane);
  stage_rows(dout + ooff, osl, L, sdo, lane);
  __syncwarp();
  dot_pairs(sq, sk, ss, L, p.scale, lane);
  dot_pairs(sdo, sv, sd, L, 1.0f, lane);
  __syncwarp();
  softmax_row(ss, L, lane);
  if (lane < L) {
    bool keep[SA_LMAX];
    const bool drop = p.drop_p > 0.f;
    if (drop) row_keep(p.seed, bh, lane, p.drop_p, keep);
    const float ks = drop ? 1.0f / (1.0f - p.drop_p) : 1.0f;
    float* rp = ss + lane * SA_LP;
    float* rd = sd + lane * SA_LP;
    float delta = 0.f;
#pragma unroll
    for (int j = 0; j < SA_LMAX; ++j) {
      if (j < L) {
        const float m = (!drop || keep[j]) ? ks : 0.f;
        rd[j] *= m;                          // dP = mask * d(dropped P) / (1 - p)
        delta = fmaf(rd[j], rp[j], delta);
      }
    }
#pragma unroll
    for (int j = 0; j < SA_LMAX; ++j) {
      if (j < L) {
        const float m = (!drop || keep[j]) ? ks : 0.f;
        const float pj = rp[j];
        rd[j] = pj * (rd[j] - delta) * p.scale;
        rp[j] = pj * m;
      }
    }
  }
  __syncwarp();
  {
    float dv0[SA_LMAX], dv1[SA_LMAX], dk0[SA_LMAX], dk1[SA_LMAX];
#pragma unroll
    for (int j = 0; j < SA_LMAX; ++j) dv0[j] = dv1[j] = dk0[j] = dk1[j] = 0.f;
    for (int i = 0; i < L; ++i) {
      const float g0 = sdo[i * SA_RS + lane], g1 = sdo[i * SA_RS + lane + 32];
      const float q0 = sq[i * SA_RS + lane], q1 = sq[i * SA_RS + lane + 32];
#pragma unroll
      for (int j = 0; j < SA_LMAX; ++j) {
        if (j < L) {
          const float pd = ss[i * SA_LP + j], ds = sd[i * SA_LP + j];
          dv0[j] = fmaf(pd, g0, dv0[j]);
          dv1[j] = fmaf(pd, g1, dv1[j]);
          dk0[j] = fmaf(ds, q0, dk0[j]);
          dk1[j] = fmaf(ds, q1, dk1[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < SA_LMAX; ++j) {
      if (j < L) {
        stf(dv + ooff + j * osl + lane, dv0[j]);
        stf(dv + ooff + j * osl + lane + 32, dv1[j]);
        stf(dk + ooff + j * osl + lane, dk0[j]);
        stf(dk + ooff + j * osl + lane + 32, dk1[j]);
      }
    }
  }
  for (int i = 0; i < L; ++i) {
    float a0 = 0.f, a1 = 0.f;
    for (int j = 0; j < L; ++j) {
      const float ds = sd[i * SA_LP + j];
      a0 = fmaf(ds, sk[j * SA_RS + lane], a0);
      a1 = fmaf(ds, sk[j * SA_RS + lane + 32], a1);
    }
    stf(dq + ooff + i * osl + lane, a0);
    stf(dq + ooff + i * osl + lane + 32, a1);
  }
}

int sa_check(const SaParams& p, int D, int dtype) {
  B2D_REQUIRE(p.q && p.k && p.v && p.B > 0 && p.H > 0 && p.L > 0, B2D_ERR_SHAPE, "short_attention: bad arguments");
  B2D_REQUIRE(D == SA_D && p.L <= SA_LMAX, B2D_ERR_SHAPE,
              "short_attention: built for head dim %d and at most %d tokens (got %d, %d)", SA_D, SA_LMAX, D, p.L);
  B2D_REQUIRE(dtype >= 0 && dtype <= 2, B2D_ERR_SHAPE, "short_attention: dtype code %d (0 fp32, 1 bf16, 2 fp16)", dtype);
  B2D_REQUIRE(p.drop_p >= 0.f && p.drop_p < 1.f, B2D_ERR_SHAPE, "short_attention: dropout probability %f", p.drop_p);
  return B2D_OK;
}

constexpr size_t SA_FWD_SMEM = SA_WARPS * (2 * SA_LMAX * SA_RS + SA_LMAX * SA_LP) * sizeof(float);
constexpr size_t SA_BWD_SMEM = SA_WARPS * (4 * SA_LMAX * SA_RS + 2 * SA_LMAX * SA_LP) * sizeof(float);

template <typename T>
int sa_launch_fwd(const SaParams& p, void* out, cudaStream_t st) {
  const unsigned grid = (unsigned)(((long long)p.B * p.H + SA_WARPS - 1) / SA_WARPS);
  short_attention_fwd_kernel<T><<<grid, SA_WARPS * 32, SA_FWD_SMEM, st>>>(p, (T*)out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

template <typename T>
int sa_launch_bwd(const SaParams& p, const void* dout, void* dq, void* dk, void* dv, cudaStream_t st) {
  B2D_CHECK_CUDA(cudaFuncSetAttribute(short_attention_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)SA_BWD_SMEM));
  const unsigned grid = (unsigned)(((long long)p.B * p.H + SA_WARPS - 1) / SA_WARPS);
  short_attention_bwd_kernel<T><<<grid, SA_WARPS * 32, SA_BWD_SMEM, st>>>(p, (const T*)dout, (T*)dq, (T*)dk, (T*)dv);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

}  // namespace
}  // namespace b2d

using namespace b2d;

extern "C" int b2d_short_attention_fwd(const void* q, const void* k, const void* v, int64_t sb, int64_t sl, int64_t sh,
                                       int B, int H, int L, int D, int dtype, float scale, float drop_p, uint64_t seed,
                                       void* out, void* stream) {
  SaParams p{q, k, v, sb, sl, sh, B, H, L, scale, drop_p, seed};
  if (int e = sa_check(p, D, dtype)) return e;
  B2D_REQUIRE(out, B2D_ERR_SHAPE, "short_attention_fwd: null output");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0) return sa_launch_fwd<float>(p, out, st);
  if (dtype == 1) return sa_launch_fwd<__nv_bfloat16>(p, out, st);
  return sa_launch_fwd<__half>(p, out, st);
}

extern "C" int b2d_short_attention_bwd(const void* q, const void* k, const void* v, int64_t sb, int64_t sl, int64_t sh,
                                       const void* dout, int B, int H, int L, int D, int dtype, float scale,
                                       float drop_p, uint64_t seed, void* dq, void* dk, void* dv, void* stream) {
  SaParams p{q, k, v, sb, sl, sh, B, H, L, scale, drop_p, seed};
  if (int e = sa_check(p, D, dtype)) return e;
  B2D_REQUIRE(dout && dq && dk && dv, B2D_ERR_SHAPE, "short_attention_bwd: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0) return sa_launch_bwd<float>(p, dout, dq, dk, dv, st);
  if (dtype == 1) return sa_launch_bwd<__nv_bfloat16>(p, dout, dq, dk, dv, st);
  return sa_launch_bwd<__half>(p, dout, dq, dk, dv, st);
}
