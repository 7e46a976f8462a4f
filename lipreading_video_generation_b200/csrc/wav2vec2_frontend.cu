// wav2vec2_frontend.cu -- the first feature-extractor layer of the audio encoder the reference instantiates in
// unet_audio.py:10-18 (HF transformers `Wav2Vec2Model`, third-party; layer class `Wav2Vec2GroupNormConvLayer`:
// Conv1d(1 -> C, kernel 10, stride 5, no padding) -> GroupNorm(C groups of one channel, affine) -> GELU(erf)).
//
// At the training shape (256 windows x 4000 samples -> (256, 512, 799)) this single layer produces the largest
// activation of the whole encoder; run as separate library ops (conv, fp32 GroupNorm with its casts, GELU and their
// backwards) it moves that 105 M-element tensor ~20 times.  Here the waveform window sits in shared memory, the conv
// output of one (window, channel) row lives in the registers of one warp, and the row is written once (forward) or
// read once (backward: d_out) -- the only HBM traffic left.
//
//   forward : out[n,c,t] = gelu(gamma_c * (y - mean_nc) * rstd_nc + beta_c),  y[n,c,t] = b_c + sum_k w[c,k] a[n, t*stride+k]
//   backward: d_w[c,k], d_b[c], d_gamma[c], d_beta[c]  (+=; the waveform needs no gradient)
#include "common.cuh"
#include <cuda_bf16.h>

namespace b2d {

constexpr int W2V_K = 10;     // kernel width of layer 0 (Wav2Vec2Config.conv_kernel[0])
constexpr int W2V_TPL = 25;   // time steps per lane: T <= 800 (4000-sample windows give 799)
constexpr float RSQRT2 = 0.70710678118654752f;
constexpr float RSQRT2PI = 0.39894228040143268f;

struct W2vParams {
  const float* audio; long long ld_audio;
  const float* w; const float* bias; const float* gamma; const float* beta;
  int N, S, C, T, stride;
  float eps;
};

template <typename T> __device__ __forceinline__ float w2v_ld(const T* p) { return (float)*p; }
template <> __device__ __forceinline__ float w2v_ld<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float w2v_ld<__half>(const __half* p) { return __half2float(*p); }
template <typename T> __device__ __forceinline__ void w2v_st(T* p, float v) { *p = v; }
template <> __device__ __forceinline__ void w2v_st<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void w2v_st<__half>(__half* p, float v) { *p = __float2half_rn(v); }

// Normal cdf / pdf at z without the data-dependent branch of erff() (whose two ranges diverge inside nearly every warp
// for N(0,1)-like z): erf by Abramowitz & Stegun 7.1.26, |error| <= 1.5e-7, sharing exp(-z^2/2) with the pdf.
// One MUFU.RCP + one MUFU.EX2 + 8 FMA-class instructions.
__device__ __forceinline__ void normal_cdf_pdf(float z, float& cdf, float& pdf) {
  const float u = fabsf(z) * RSQRT2;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, u, 1.0f)));
  const float e = fast_exp2(-0.72134752044448170f * z * z);   // exp(-z^2 / 2)
  float q = fmaf(1.061405429f, t, -1.453152027f);
  q = fmaf(q, t, 1.421413741f);
  q = fmaf(q, t, -0.284496736f);
  q = fmaf(q, t, 0.254829592f);
  q = 0.5f * q * t * e;
  cdf = z >= 0.f ? 1.0f - q : q;
  pdf = e * RSQRT2PI;
}

__device__ __forceinline__ void load_window(float* sa, const W2vParams& p, int n) {
  const float* a = p.audio + (long long)n * p.ld_audio;
  for (int i = threadIdx.x; i < p.S; i += blockDim.x) sa[i] = __ldg(a + i);
}

// grid = (ceil(C / CPB), N), 256 threads; a warp owns channels warp, warp + 8, ... of the block's CPB channels.
constexpr int W2V_FWD_CPB = 32;

template <typename OutT>
__global__ void __launch_bounds__(256)
w2v_layer0_fwd_kernel(W2vParams p, OutT* __restrict__ out, float* __restrict__ mean, float* __restrict__ rstd) {
  extern __shared__ float sa[];
  const int n = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  load_window(sa, p, n);
  __syncthreads();
  const float inv_T = 1.0f / (float)p.T;
  for (int cc = warp; cc < W2V_FWD_CPB; cc += 8) {
    const int c = blockIdx.x * W2V_FWD_CPB + cc;
    if (c >= p.C) break;
    float w[W2V_K];
#pragma unroll
    for (int k = 0; k < W2V_K; ++k) w[k] = __ldg(p.w + c * W2V_K + k);
    const float b = p.bias ? __ldg(p.bias + c) : 0.f;
    float y[W2V_TPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < W2V_TPL; ++i) {
      const int t = lane + 32 * i;
      float acc = 0.f;
      if (t < p.T) {
        acc = b;
        const float* a = sa + t * p.stride;
#pragma unroll
        for (int k = 0; k < W2V_K; ++k) acc = fmaf(w[k], a[k], acc);
        s += acc;
      }
      y[i] = acc;
    }
    const float mu = warp_sum(s) * inv_T;
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < W2V_TPL; ++i) {
      const float d = y[i] - mu;
      if (lane + 32 * i < p.T) v = fmaf(d, d, v);
    }
    const float rs = rsqrtf(warp_sum(v) * inv_T + p.eps);
    const float g = __ldg(p.gamma + c) * rs;
    const float bb = __ldg(p.beta + c);
    OutT* o = out + ((long long)n * p.C + c) * p.T;
#pragma unroll
    for (int i = 0; i < W2V_TPL; ++i) {
      const int t = lane + 32 * i;
      const float z = fmaf(y[i] - mu, g, bb);   // centred first: well conditioned even for a constant row
      float cdf, pdf;
      normal_cdf_pdf(z, cdf, pdf);
      if (t < p.T) w2v_st(o + t, z * cdf);
    }
    if (lane == 0) {
      mean[(long long)n * p.C + c] = mu;
      rstd[(long long)n * p.C + c] = rs;
    }
  }
}

// grid = (ceil(C / 8), ceil(N / n_chunk)), 256 threads; a warp owns one channel and walks the windows of its chunk,
// keeping the parameter-gradient partial sums in registers; one set of atomics per (channel, chunk).
template <typename GT>
__global__ void __launch_bounds__(256, 2)
w2v_layer0_bwd_kernel(W2vParams p, const GT* __restrict__ dout, const float* __restrict__ mean,
                      const float* __restrict__ rstd, float* __restrict__ dw, float* __restrict__ dbias,
                      float* __restrict__ dgamma, float* __restrict__ dbeta, int n_chunk) {
  extern __shared__ float sa[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = blockIdx.x * 8 + warp;
  const bool active = c < p.C;
  const int n0 = blockIdx.y * n_chunk;
  const int n1 = min(p.N, n0 + n_chunk);
  float w[W2V_K], aw[W2V_K];
#pragma unroll
  for (int k = 0; k < W2V_K; ++k) { w[k] = active ? __ldg(p.w + c * W2V_K + k) : 0.f; aw[k] = 0.f; }
  const float b = (active && p.bias) ? __ldg(p.bias + c) : 0.f;
  const float gam = active ? __ldg(p.gamma + c) : 0.f, bet = active ? __ldg(p.beta + c) : 0.f;
  const float inv_T = 1.0f / (float)p.T;
  float a_gamma = 0.f, a_beta = 0.f, a_bias = 0.f;
  for (int n = n0; n < n1; ++n) {
    __syncthreads();
    load_window(sa, p, n);
    __syncthreads();
    if (!active) continue;
    const float mu = __ldg(mean + (long long)n * p.C + c), rs = __ldg(rstd + (long long)n * p.C + c);
    const GT* g_in = dout + ((long long)n * p.C + c) * p.T;
    float gz[W2V_TPL];   // d_out * gelu'(z); x_hat is recomputed in the second pass from the taps it loads anyway
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < W2V_TPL; ++i)      // all of the row's d_out loads in flight before the first use
      gz[i] = (lane + 32 * i < p.T) ? w2v_ld(g_in + lane + 32 * i) : 0.f;
#pragma unroll
    for (int i = 0; i < W2V_TPL; ++i) {
      const int t = lane + 32 * i;
      float dz = 0.f;
      if (t < p.T) {
        float acc = b;
        const float* a = sa + t * p.stride;
#pragma unroll
        for (int k = 0; k < W2V_K; ++k) acc = fmaf(w[k], a[k], acc);
        const float x = (acc - mu) * rs;
        const float z = fmaf(x, gam, bet);
        float cdf, pdf;
        normal_cdf_pdf(z, cdf, pdf);
        dz = gz[i] * fmaf(z, pdf, cdf);
        s1 += dz;
        s2 = fmaf(dz, x, s2);
      }
      gz[i] = dz;
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    a_beta += s1;
    a_gamma += s2;
    const float m1 = gam * s1 * inv_T, m2 = gam * s2 * inv_T;
#pragma unroll
    for (int i = 0; i < W2V_TPL; ++i) {
      const int t = lane + 32 * i;
      if (t < p.T) {
        const float* a = sa + t * p.stride;
        float av[W2V_K];
        float acc = b;
#pragma unroll
        for (int k = 0; k < W2V_K; ++k) { av[k] = a[k]; acc = fmaf(w[k], av[k], acc); }
        const float x = (acc - mu) * rs;
        const float dy = rs * (fmaf(gam, gz[i], -m1) - x * m2);
        a_bias += dy;
#pragma unroll
        for (int k = 0; k < W2V_K; ++k) aw[k] = fmaf(dy, av[k], aw[k]);
      }
    }
  }
  if (!active) return;
#pragma unroll
  for (int k = 0; k < W2V_K; ++k) aw[k] = warp_sum(aw[k]);
  a_bias = warp_sum(a_bias);
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < W2V_K; ++k) atomicAdd(dw + c * W2V_K + k, aw[k]);
    atomicAdd(dgamma + c, a_gamma);   // a_gamma / a_beta are already warp-uniform totals
    atomicAdd(dbeta + c, a_beta);
    if (dbias) atomicAdd(dbias + c, a_bias);
  }
}

static int w2v_check(const W2vParams& p, int K) {
  B2D_REQUIRE(p.audio && p.w && p.gamma && p.beta && p.N > 0 && p.C > 0 && p.S > 0, B2D_ERR_SHAPE,
              "w2v_layer0: bad arguments");
  B2D_REQUIRE(K == W2V_K && p.stride > 0 && p.S >= K, B2D_ERR_SHAPE,
              "w2v_layer0: kernel width %d / stride %d / %d samples unsupported (built for width %d)", K, p.stride, p.S,
              W2V_K);
  B2D_REQUIRE(p.T <= 32 * W2V_TPL, B2D_ERR_SHAPE, "w2v_layer0: %d output steps exceed the %d a warp keeps in registers",
              p.T, 32 * W2V_TPL);
  B2D_REQUIRE((size_t)p.S * sizeof(float) <= 200 * 1024, B2D_ERR_SHAPE, "w2v_layer0: window of %d samples exceeds shared memory", p.S);
  return B2D_OK;
}

template <typename Kern> static int w2v_smem(Kern k, size_t bytes) {
  if (bytes > 48 * 1024) B2D_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return B2D_OK;
}

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_w2v_layer0_fwd(const float* audio, int64_t ld_audio, const float* w, const float* bias,
                                  const float* gamma, const float* beta, int N, int S, int C, int K, int stride,
                                  float eps, void* out, int out_dtype, float* mean, float* rstd, void* stream) {
  W2vParams p{audio, ld_audio, w, bias, gamma, beta, N, S, C, S >= K && stride > 0 ? (S - K) / stride + 1 : 0, stride, eps};
  if (int e = w2v_check(p, K)) return e;
  B2D_REQUIRE(out && mean && rstd && out_dtype >= 0 && out_dtype <= 2, B2D_ERR_SHAPE, "w2v_layer0_fwd: bad outputs");
  const dim3 grid((C + W2V_FWD_CPB - 1) / W2V_FWD_CPB, N);
  const size_t smem = (size_t)S * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == 0) {
    if (int e = w2v_smem(w2v_layer0_fwd_kernel<float>, smem)) return e;
    w2v_layer0_fwd_kernel<float><<<grid, 256, smem, st>>>(p, (float*)out, mean, rstd);
  } else if (out_dtype == 1) {
    if (int e = w2v_smem(w2v_layer0_fwd_kernel<__nv_bfloat16>, smem)) return e;
    w2v_layer0_fwd_kernel<__nv_bfloat16><<<grid, 256, smem, st>>>(p, (__nv_bfloat16*)out, mean, rstd);
  } else {
    if (int e = w2v_smem(w2v_layer0_fwd_kernel<__half>, smem)) return e;
    w2v_layer0_fwd_kernel<__half><<<grid, 256, smem, st>>>(p, (__half*)out, mean, rstd);
  }
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_w2v_layer0_bwd(const float* audio, int64_t ld_audio, const float* w, const float* bias,
                                  const float* gamma, const float* beta, int N, int S, int C, int K, int stride,
                                  float eps, const void* dout, int dout_dtype, const float* mean, const float* rstd,
                                  float* dw, float* dbias, float* dgamma, float* dbeta, void* stream) {
  W2vParams p{audio, ld_audio, w, bias, gamma, beta, N, S, C, S >= K && stride > 0 ? (S - K) / stride + 1 : 0, stride, eps};
  if (int e = w2v_check(p, K)) return e;
  B2D_REQUIRE(dout && mean && rstd && dw && dgamma && dbeta && dout_dtype >= 0 && dout_dtype <= 2, B2D_ERR_SHAPE,
              "w2v_layer0_bwd: bad arguments");
  const int n_chunk = N >= 64 ? 8 : 1;
  const dim3 grid((C + 7) / 8, (N + n_chunk - 1) / n_chunk);
  const size_t smem = (size_t)S * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  if (dout_dtype == 0) {
    if (int e = w2v_smem(w2v_layer0_bwd_kernel<float>, smem)) return e;
    w2v_layer0_bwd_kernel<float><<<grid, 256, smem, st>>>(p, (const float*)dout, mean, rstd, dw, dbias, dgamma, dbeta, n_chunk);
  } else if (dout_dtype == 1) {
    if (int e = w2v_smem(w2v_layer0_bwd_kernel<__nv_bfloat16>, smem)) return e;
    w2v_layer0_bwd_kernel<__nv_bfloat16><<<grid, 256, smem, st>>>(p, (const __nv_bfloat16*)dout, mean, rstd, dw, dbias, dgamma, dbeta, n_chunk);
  } else {
    if (int e = w2v_smem(w2v_layer0_bwd_kernel<__half>, smem)) return e;
    w2v_layer0_bwd_kernel<__half><<<grid, 256, smem, st>>>(p, (const __half*)dout, mean, rstd, dw, dbias, dgamma, dbeta, n_chunk);
  }
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// LayerNorm over the last dimension for the encoder's 26 nn.LayerNorm calls per forward (HF `Wav2Vec2EncoderLayer`,
// `Wav2Vec2Encoder`, `Wav2Vec2FeatureProjection`; third-party arithmetic, row N4): 3072 rows x 768 at the train64 batch.
// torch serves the backward with `layer_norm_grad_input_kernel` + `GammaBetaBackwardCUDAKernelTemplate` -- 9 + 30 us per
// call for 9 MB of traffic (profiles/r2_launch_list.md) -- because its two-stage gamma / beta reduction is built for
// many more rows.  Here one warp owns a row (C / 32 values per lane in registers), walks several rows and keeps its
// gamma / beta partial sums in registers across them; one shared-memory reduction per CTA, one atomic per column per CTA.
namespace b2d {
namespace {

constexpr int LN_WARPS = 8;
constexpr int LN_MAX_PER_LANE = 32;     // C <= 1024

template <typename T> __device__ __forceinline__ float ln_ld(const T* p) { return (float)*p; }
template <> __device__ __forceinline__ float ln_ld<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float ln_ld<__half>(const __half* p) { return __half2float(*p); }

__device__ __forceinline__ float ln_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T, int CPL>   // CPL = C / 32
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_fwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta, int R,
                     float eps, float* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd) {
  constexpr int C = CPL * 32;
  const int lane = threadIdx.x & 31;
  const int wid = blockIdx.x * LN_WARPS + (threadIdx.x >> 5), nw = gridDim.x * LN_WARPS;
  float g[CPL], b[CPL];
#pragma unroll
  for (int i = 0; i < CPL; ++i) { g[i] = __ldg(gamma + lane + 32 * i); b[i] = __ldg(beta + lane + 32 * i); }
  for (int r = wid; r < R; r += nw) {
    float v[CPL];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < CPL; ++i) { v[i] = ln_ld(x + (long long)r * C + lane + 32 * i); s += v[i]; }
    const float mu = ln_warp_sum(s) * (1.0f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < CPL; ++i) { const float d = v[i] - mu; q = fmaf(d, d, q); }
    const float rs = rsqrtf(ln_warp_sum(q) * (1.0f / C) + eps);
#pragma unroll
    for (int i = 0; i < CPL; ++i) y[(long long)r * C + lane + 32 * i] = fmaf((v[i] - mu) * rs, g[i], b[i]);
    if (lane == 0) { mean[r] = mu; rstd[r] = rs; }
  }
}

template <typename T, int CPL>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_bwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const float* __restrict__ dy, int R, float* __restrict__ dx,
                     float* __restrict__ dgamma, float* __restrict__ dbeta) {
  constexpr int C = CPL * 32;
  __shared__ float sh[2][C];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wid = blockIdx.x * LN_WARPS + warp, nw = gridDim.x * LN_WARPS;
  for (int i = threadIdx.x; i < 2 * C; i += LN_WARPS * 32) (&sh[0][0])[i] = 0.f;
  __syncthreads();
  float g[CPL], ag[CPL], ab[CPL];
#pragma unroll
  for (int i = 0; i < CPL; ++i) { g[i] = __ldg(gamma + lane + 32 * i); ag[i] = 0.f; ab[i] = 0.f; }
  for (int r = wid; r < R; r += nw) {
    const float mu = __ldg(mean + r), rs = __ldg(rstd + r);
    float xh[CPL], gy[CPL];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
      const long long o = (long long)r * C + lane + 32 * i;
      const float d = __ldg(dy + o);
      xh[i] = (ln_ld(x + o) - mu) * rs;
      gy[i] = d * g[i];
      s1 += gy[i];
      s2 = fmaf(gy[i], xh[i], s2);
      ab[i] += d;
      ag[i] = fmaf(d, xh[i], ag[i]);
    }
    s1 = ln_warp_sum(s1) * (1.0f / C);
    s2 = ln_warp_sum(s2) * (1.0f / C);
#pragma unroll
    for (int i = 0; i < CPL; ++i) dx[(long long)r * C + lane + 32 * i] = rs * (gy[i] - s1 - xh[i] * s2);
  }
#pragma unroll
  for (int i = 0; i < CPL; ++i) { atomicAdd(&sh[0][lane + 32 * i], ag[i]); atomicAdd(&sh[1][lane + 32 * i], ab[i]); }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += LN_WARPS * 32) { atomicAdd(dgamma + i, sh[0][i]); atomicAdd(dbeta + i, sh[1][i]); }
}

template <typename T>
int launch_ln_fwd(int cpl, int blocks, cudaStream_t st, const void* x, const float* g, const float* b, int R, float eps,
                  float* y, float* mean, float* rstd) {
  switch (cpl) {
#define B2D_LN_CASE(N) case N: layernorm_fwd_kernel<T, N><<<blocks, LN_WARPS * 32, 0, st>>>((const T*)x, g, b, R, eps, y, mean, rstd); break;
    B2D_LN_CASE(2) B2D_LN_CASE(4) B2D_LN_CASE(8) B2D_LN_CASE(16) B2D_LN_CASE(24) B2D_LN_CASE(32)
#undef B2D_LN_CASE
    default: return 1;
  }
  return 0;
}
template <typename T>
int launch_ln_bwd(int cpl, int blocks, cudaStream_t st, const void* x, const float* g, const float* mean, const float* rstd,
                  const float* dy, int R, float* dx, float* dg, float* db) {
  switch (cpl) {
#define B2D_LN_CASE(N) case N: layernorm_bwd_kernel<T, N><<<blocks, LN_WARPS * 32, 0, st>>>((const T*)x, g, mean, rstd, dy, R, dx, dg, db); break;
    B2D_LN_CASE(2) B2D_LN_CASE(4) B2D_LN_CASE(8) B2D_LN_CASE(16) B2D_LN_CASE(24) B2D_LN_CASE(32)
#undef B2D_LN_CASE
    default: return 1;
  }
  return 0;
}

int ln_blocks(int R) {
  int b = (R + LN_WARPS * 4 - 1) / (LN_WARPS * 4);      // ~4 rows per warp: the gamma / beta partial sums amortise
  const int cap = 2 * num_sms();
  if (b > cap) b = cap;
  return b < 1 ? 1 : b;
}

}  // namespace
}  // namespace b2d

extern "C" int b2d_layernorm_fwd(const void* x, int x_dtype, const float* gamma, const float* beta, int R, int C, float eps,
                                 float* y, float* mean, float* rstd, void* stream) {
  B2D_REQUIRE(x && gamma && beta && y && mean && rstd && R > 0, B2D_ERR_SHAPE, "layernorm_fwd: bad arguments");
  B2D_REQUIRE(C % 32 == 0 && C <= 32 * b2d::LN_MAX_PER_LANE, B2D_ERR_SHAPE, "layernorm_fwd: C=%d", C);
  const int blocks = b2d::ln_blocks(R);
  cudaStream_t st = (cudaStream_t)stream;
  int bad;
  if (x_dtype == 0)      bad = b2d::launch_ln_fwd<float>(C / 32, blocks, st, x, gamma, beta, R, eps, y, mean, rstd);
  else if (x_dtype == 1) bad = b2d::launch_ln_fwd<__nv_bfloat16>(C / 32, blocks, st, x, gamma, beta, R, eps, y, mean, rstd);
  else if (x_dtype == 2) bad = b2d::launch_ln_fwd<__half>(C / 32, blocks, st, x, gamma, beta, R, eps, y, mean, rstd);
  else bad = 1;
  B2D_REQUIRE(!bad, B2D_ERR_SHAPE, "layernorm_fwd: unsupported C=%d (C / 32 in {2,4,8,16,24,32}) or dtype %d", C, x_dtype);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_layernorm_bwd(const void* x, int x_dtype, const float* gamma, const float* mean, const float* rstd,
                                 const float* dy, int R, int C, float* dx, float* dgamma, float* dbeta, void* stream) {
  B2D_REQUIRE(x && gamma && mean && rstd && dy && dx && dgamma && dbeta && R > 0, B2D_ERR_SHAPE, "layernorm_bwd: bad arguments");
  B2D_REQUIRE(C % 32 == 0 && C <= 32 * b2d::LN_MAX_PER_LANE, B2D_ERR_SHAPE, "layernorm_bwd: C=%d", C);
  const int blocks = b2d::ln_blocks(R);
  cudaStream_t st = (cudaStream_t)stream;
  int bad;
  if (x_dtype == 0)      bad = b2d::launch_ln_bwd<float>(C / 32, blocks, st, x, gamma, mean, rstd, dy, R, dx, dgamma, dbeta);
  else if (x_dtype == 1) bad = b2d::launch_ln_bwd<__nv_bfloat16>(C / 32, blocks, st, x, gamma, mean, rstd, dy, R, dx, dgamma, dbeta);
  else if (x_dtype == 2) bad = b2d::launch_ln_bwd<__half>(C / 32, blocks, st, x, gamma, mean, rstd, dy, R, dx, dgamma, dbeta);
  else bad = 1;
  B2D_REQUIRE(!bad, B2D_ERR_SHAPE, "layernorm_bwd: unsupported C=%d (C / 32 in {2,4,8,16,24,32}) or dtype %d", C, x_dtype);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Bias gradient of the encoder's 73 nn.Linear layers: out[c] = sum_r dy[r][c] over a 16-bit (R, C) gradient.  torch
// serves it with `reduce_kernel<sum, bf16>` (15 us at 3072 x 768) plus a cast of the result to fp32 (4.6 us) -- twice the
// time of the layer's dW GEMM (profiles/r2_launch_list.md).  CTA = 64 columns x a slice of rows; 8 row lanes per column
// pair, fp32 accumulation, one atomic per column per CTA into the zeroed fp32 output.
namespace b2d {
namespace {

template <typename T2>   // __nv_bfloat162 or __half2
__device__ __forceinline__ float2 cs_ld2(const void* p);
template <> __device__ __forceinline__ float2 cs_ld2<__nv_bfloat162>(const void* p) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
}
template <> __device__ __forceinline__ float2 cs_ld2<__half2>(const void* p) {
  return __half22float2(*reinterpret_cast<const __half2*>(p));
}

template <typename T2>
__global__ void __launch_bounds__(256)
colsum_rows16_kernel(const uint16_t* __restrict__ x, long long ld, int R, int C, int rows_per_cta, float* __restrict__ out) {
  __shared__ float sh[8][65];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 64 + tx * 2;
  const int r_lo = blockIdx.y * rows_per_cta, r_hi = min(R, r_lo + rows_per_cta);
  float a0 = 0.f, a1 = 0.f;
  if (c < C) {
    for (int r = r_lo + ty; r < r_hi; r += 32) {
      float2 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        v[u] = (r + 8 * u < r_hi) ? cs_ld2<T2>(x + (long long)(r + 8 * u) * ld + c) : make_float2(0.f, 0.f);
      a0 += (v[0].x + v[1].x) + (v[2].x + v[3].x);
      a1 += (v[0].y + v[1].y) + (v[2].y + v[3].y);
    }
  }
  sh[ty][tx * 2] = a0; sh[ty][tx * 2 + 1] = a1;
  __syncthreads();
  if (threadIdx.x < 64 && blockIdx.x * 64 + (int)threadIdx.x < C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sh[i][threadIdx.x];
    atomicAdd(out + blockIdx.x * 64 + threadIdx.x, t);
  }
}

}  // namespace
}  // namespace b2d

extern "C" int b2d_colsum_rows16(const void* x, int x_dtype, int64_t ld, int R, int C, float* out, void* stream) {
  B2D_REQUIRE(x && out && R > 0 && C > 0 && C % 2 == 0 && ld % 2 == 0 && (reinterpret_cast<uintptr_t>(x) & 3) == 0,
              B2D_ERR_SHAPE, "colsum_rows16: bad arguments (C=%d, ld=%lld must be even, x 4-byte aligned)", C, (long long)ld);
  B2D_REQUIRE(x_dtype == 1 || x_dtype == 2, B2D_ERR_SHAPE, "colsum_rows16: dtype %d (1 = bf16, 2 = fp16)", x_dtype);
  cudaStream_t st = (cudaStream_t)stream;
  B2D_CHECK_CUDA(cudaMemsetAsync(out, 0, (size_t)C * sizeof(float), st));
  const int gx = (C + 63) / 64;
  int gy = (2 * b2d::num_sms() + gx - 1) / gx;
  if (gy > (R + 31) / 32) gy = (R + 31) / 32;
  if (gy < 1) gy = 1;
  const int rows_per_cta = (R + gy - 1) / gy;
  dim3 grid((unsigned)gx, (unsigned)gy);
  if (x_dtype == 1) b2d::colsum_rows16_kernel<__nv_bfloat162><<<grid, 256, 0, st>>>((const uint16_t*)x, ld, R, C, rows_per_cta, out);
  else              b2d::colsum_rows16_kernel<__half2><<<grid, 256, 0, st>>>((const uint16_t*)x, ld, R, C, rows_per_cta, out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// (B, R, C) -> (B, C, R) transpose of 16-bit elements through 64 x 64 shared-memory tiles: the one layout change between
// the fused first feature-extractor layer, which writes (N, 512, 799) time-contiguous rows, and the channels-last
// tensor the library convolutions of layers 1..6 run on (and back for the gradient).  torch's strided copy does this at
// ~1 TB/s (0.36-0.42 ms per direction for 209 MB in + out); tiled, both sides move in 128-byte segments.
namespace b2d {
namespace {

__global__ void __launch_bounds__(256)
transpose16_kernel(const uint16_t* __restrict__ in, uint16_t* __restrict__ out, int R, int C) {
  __shared__ uint16_t tile[64][66];
  const long long base = (long long)blockIdx.z * R * C;
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 column pairs x 8 row lanes
  for (int r = ty; r < 64; r += 8) {
    const int rr = r0 + r, cc = c0 + 2 * tx;
    uint32_t v = 0;
    if (rr < R) {
      if (cc + 1 < C && ((C & 1) == 0)) v = *reinterpret_cast<const uint32_t*>(in + base + (long long)rr * C + cc);
      else {
        if (cc < C) v = in[base + (long long)rr * C + cc];
        if (cc + 1 < C) v |= (uint32_t)in[base + (long long)rr * C + cc + 1] << 16;
      }
    }
    tile[r][2 * tx] = (uint16_t)(v & 0xffffu);
    tile[r][2 * tx + 1] = (uint16_t)(v >> 16);
  }
  __syncthreads();
  for (int c = ty; c < 64; c += 8) {
    const int cc = c0 + c, rr = r0 + 2 * tx;
    if (cc < C) {
      const uint32_t v = (uint32_t)tile[2 * tx][c] | ((uint32_t)tile[2 * tx + 1][c] << 16);
      if (rr + 1 < R && ((R & 1) == 0)) *reinterpret_cast<uint32_t*>(out + base + (long long)cc * R + rr) = v;
      else {
        if (rr < R) out[base + (long long)cc * R + rr] = (uint16_t)(v & 0xffffu);
        if (rr + 1 < R) out[base + (long long)cc * R + rr + 1] = (uint16_t)(v >> 16);
      }
    }
  }
}

}  // namespace
}  // namespace b2d

extern "C" int b2d_transpose16(const void* in, void* out, int B, int R, int C, void* stream) {
  B2D_REQUIRE(in && out && B > 0 && R > 0 && C > 0 && B <= 65535, B2D_ERR_SHAPE, "transpose16: bad arguments");
  B2D_REQUIRE((reinterpret_cast<uintptr_t>(in) & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0, B2D_ERR_ALIGN,
              "transpose16: buffers must be 4-byte aligned");
  dim3 grid((unsigned)((C + 63) / 64), (unsigned)((R + 63) / 64), (unsigned)B);
  B2D_REQUIRE(grid.y <= 65535, B2D_ERR_SHAPE, "transpose16: R=%d too large", R);
  b2d::transpose16_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)in, (uint16_t*)out, R, C);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
