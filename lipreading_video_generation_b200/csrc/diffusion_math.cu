// diffusion_math.cu -- q_sample, sampler updates, MSE loss (fp32, HBM bound).
// Reference semantics: video-generation/diffusion/linear_noise_scheduler.py:24-101,
// noise_scheduler.py:13-29, train.py:103,132.  All arithmetic uses explicit
// round-to-nearest intrinsics in the reference's operation order so ptxas cannot
// contract mul+add into FMA: results are bit-identical to the eager PyTorch ops.
#include "common.cuh"

namespace b2d {

// ---------------------------------------------------------------- q_sample
// One CTA-row of work per (sample, chunk): grid.y = B so the schedule gather
// happens once per CTA (two scalar loads) and the body is pure 128-bit streaming.
template <bool VEC>
__global__ void __launch_bounds__(256)
q_sample_kernel(const float* __restrict__ x0, const float* __restrict__ eps,
                const int64_t* __restrict__ t, const float* __restrict__ tab_a,
                const float* __restrict__ tab_b, int T, float* __restrict__ xt,
                int64_t per_sample, int* __restrict__ err_flag) {
  const int64_t b = blockIdx.y;
  const int64_t tb = t[b];
  float a, s;
  if (tb < 0 || tb >= T) {  // reference raises IndexError; we flag + poison
    if (threadIdx.x == 0 && blockIdx.x == 0 && err_flag) atomicExch(err_flag, 1);
    a = s = __int_as_float(0x7fc00000);
  } else {
    a = __ldg(tab_a + tb);
    s = __ldg(tab_b + tb);
  }
  const float* px = x0 + b * per_sample;
  const float* pe = eps + b * per_sample;
  float* po = xt + b * per_sample;
  if (VEC) {
    const int64_t nv = per_sample >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // 2x unrolled: four 128-bit loads in flight per thread
    for (; i + stride < nv; i += 2 * stride) {
      float4 xa = ldg_stream_f4(px + 4 * i), ea = ldg_stream_f4(pe + 4 * i);
      float4 xb = ldg_stream_f4(px + 4 * (i + stride)), eb = ldg_stream_f4(pe + 4 * (i + stride));
      float4 oa, ob;
      oa.x = __fadd_rn(__fmul_rn(a, xa.x), __fmul_rn(s, ea.x));
      oa.y = __fadd_rn(__fmul_rn(a, xa.y), __fmul_rn(s, ea.y));
      oa.z = __fadd_rn(__fmul_rn(a, xa.z), __fmul_rn(s, ea.z));
      oa.w = __fadd_rn(__fmul_rn(a, xa.w), __fmul_rn(s, ea.w));
      ob.x = __fadd_rn(__fmul_rn(a, xb.x), __fmul_rn(s, eb.x));
      ob.y = __fadd_rn(__fmul_rn(a, xb.y), __fmul_rn(s, eb.y));
      ob.z = __fadd_rn(__fmul_rn(a, xb.z), __fmul_rn(s, eb.z));
      ob.w = __fadd_rn(__fmul_rn(a, xb.w), __fmul_rn(s, eb.w));
      stg_stream_f4(po + 4 * i, oa);
      stg_stream_f4(po + 4 * (i + stride), ob);
    }
    for (; i < nv; i += stride) {
      float4 xa = ldg_stream_f4(px + 4 * i), ea = ldg_stream_f4(pe + 4 * i);
      float4 oa;
      oa.x = __fadd_rn(__fmul_rn(a, xa.x), __fmul_rn(s, ea.x));
      oa.y = __fadd_rn(__fmul_rn(a, xa.y), __fmul_rn(s, ea.y));
      oa.z = __fadd_rn(__fmul_rn(a, xa.z), __fmul_rn(s, ea.z));
      oa.w = __fadd_rn(__fmul_rn(a, xa.w), __fmul_rn(s, ea.w));
      stg_stream_f4(po + 4 * i, oa);
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_sample;
         i += (int64_t)gridDim.x * blockDim.x)
      po[i] = __fadd_rn(__fmul_rn(a, px[i]), __fmul_rn(s, pe[i]));
  }
}

// ---------------------------------------------------------------- samplers
struct StepCoef {
  float s1m, sqrt_acp, beta, sqrt_alpha, sigma, sqrt_acp_prev, s1m_prev;
  int use_noise;
};
enum { STEP_DDPM = 0, STEP_V2 = 1, STEP_COS = 2, STEP_DDIM = 3 };

__device__ __forceinline__ float clamp1(float v) {
  // torch.clamp propagates NaN; fminf/fmaxf would drop it.
  return (v != v) ? v : fminf(fmaxf(v, -1.0f), 1.0f);
}

template <int KIND>
__device__ __forceinline__ void step_one(float xt, float e, float z, const StepCoef& c,
                                         float& o_prev, float& o_x0) {
  if (KIND == STEP_DDPM) {
    float x0 = __fdiv_rn(__fsub_rn(xt, __fmul_rn(c.s1m, e)), c.sqrt_acp);
    o_x0 = clamp1(x0);
    float m = __fsub_rn(xt, __fdiv_rn(__fmul_rn(c.beta, e), c.s1m));
    m = __fdiv_rn(m, c.sqrt_alpha);
    o_prev = c.use_noise ? __fadd_rn(m, __fmul_rn(c.sigma, z)) : m;
  } else if (KIND == STEP_V2) {
    float se = __fmul_rn(c.s1m, e);
    float m = __fsub_rn(xt, __fdiv_rn(se, c.sqrt_alpha));
    o_prev = __fadd_rn(m, __fmul_rn(c.sigma, z));
    o_x0 = clamp1(__fdiv_rn(__fsub_rn(xt, se), c.sqrt_acp));
  } else if (KIND == STEP_COS) {
    float m = __fdiv_rn(__fsub_rn(xt, __fmul_rn(c.s1m, e)), c.sqrt_acp);
    o_x0 = m;
    o_prev = c.use_noise ? __fadd_rn(m, __fmul_rn(c.sigma, z)) : m;
  } else {  // DDIM eta = 0
    float x0 = clamp1(__fdiv_rn(__fsub_rn(xt, __fmul_rn(c.s1m, e)), c.sqrt_acp));
    o_x0 = x0;
    o_prev = __fadd_rn(__fmul_rn(c.sqrt_acp_prev, x0), __fmul_rn(c.s1m_prev, e));
  }
}

template <int KIND, bool VEC>
__global__ void __launch_bounds__(256)
sampler_step_kernel(const float* __restrict__ xt, const float* __restrict__ eps,
                    const float* __restrict__ z, StepCoef c, float* __restrict__ x_prev,
                    float* __restrict__ x0, int64_t n) {
  const bool need_z = (KIND == STEP_V2) || ((KIND == STEP_DDPM || KIND == STEP_COS) && c.use_noise);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (VEC) {
    const int64_t nv = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
      float4 a = ldg_stream_f4(xt + 4 * i), e = ldg_stream_f4(eps + 4 * i);
      float4 zz = need_z ? ldg_stream_f4(z + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 p, q;
      step_one<KIND>(a.x, e.x, zz.x, c, p.x, q.x);
      step_one<KIND>(a.y, e.y, zz.y, c, p.y, q.y);
      step_one<KIND>(a.z, e.z, zz.z, c, p.z, q.z);
      step_one<KIND>(a.w, e.w, zz.w, c, p.w, q.w);
      stg_stream_f4(x_prev + 4 * i, p);
      stg_stream_f4(x0 + 4 * i, q);
    }
    // tail (n % 4)
    for (int64_t i = (nv << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      float p, q;
      step_one<KIND>(xt[i], eps[i], need_z ? z[i] : 0.f, c, p, q);
      x_prev[i] = p; x0[i] = q;
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      float p, q;
      step_one<KIND>(xt[i], eps[i], need_z ? z[i] : 0.f, c, p, q);
      x_prev[i] = p; x0[i] = q;
    }
  }
}

template <int KIND>
static int launch_step(const float* xt, const float* eps, const float* z, const StepCoef& c,
                       float* x_prev, float* x0, int64_t n, void* stream) {
  B2D_REQUIRE(xt && eps && x_prev && x0, B2D_ERR_SHAPE, "sampler step: null pointer");
  B2D_REQUIRE(n >= 0, B2D_ERR_SHAPE, "sampler step: negative n");
  const bool need_z = (KIND == STEP_V2) || ((KIND == STEP_DDPM || KIND == STEP_COS) && c.use_noise);
  B2D_REQUIRE(!need_z || z, B2D_ERR_SHAPE, "sampler step: z is required for this step");
  if (n == 0) return B2D_OK;
  const bool vec = aligned16(xt) && aligned16(eps) && aligned16(x_prev) && aligned16(x0) &&
                   (!need_z || aligned16(z));
  const int threads = 256;
  int64_t work = vec ? (n + 3) / 4 : n;
  int64_t blocks = (work + threads - 1) / threads;
  int64_t cap = (int64_t)num_sms() * 8;  // 8 resident CTAs of 256 threads per SM
  if (blocks > cap) blocks = cap;
  cudaStream_t s = (cudaStream_t)stream;
  if (vec) sampler_step_kernel<KIND, true><<<(unsigned)blocks, threads, 0, s>>>(xt, eps, z, c, x_prev, x0, n);
  else     sampler_step_kernel<KIND, false><<<(unsigned)blocks, threads, 0, s>>>(xt, eps, z, c, x_prev, x0, n);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

// ---------------------------------------------------------------- MSE
constexpr int MSE_THREADS = 512;

__device__ __forceinline__ float block_sum_512(float v, float* sh) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x < 32) {
    r = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.f;
    r = warp_sum(r);
  }
  return r;  // valid in warp 0
}

__global__ void __launch_bounds__(MSE_THREADS)
mse_fwd_kernel(const float* __restrict__ pred, const float* __restrict__ target, int64_t n,
               float* __restrict__ partial, unsigned int* __restrict__ counter,
               float* __restrict__ loss, int vec) {
  __shared__ float sh[32];
  __shared__ bool is_last;
  float acc0 = 0.f, acc1 = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const int64_t nv = n >> 2;
    for (int64_t i = tid; i < nv; i += stride) {
      float4 p = ldg_stream_f4(pred + 4 * i), q = ldg_stream_f4(target + 4 * i);
      float dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z, dw = p.w - q.w;
      acc0 = fmaf(dx, dx, acc0); acc1 = fmaf(dy, dy, acc1);
      acc0 = fmaf(dz, dz, acc0); acc1 = fmaf(dw, dw, acc1);
    }
    for (int64_t i = (nv << 2) + tid; i < n; i += stride) { float d = pred[i] - target[i]; acc0 = fmaf(d, d, acc0); }
  } else {
    for (int64_t i = tid; i < n; i += stride) { float d = pred[i] - target[i]; acc0 = fmaf(d, d, acc0); }
  }
  float r = block_sum_512(acc0 + acc1, sh);
  if (threadIdx.x == 0) {
    partial[blockIdx.x] = r;
    __threadfence();
    unsigned int done = atomicAdd(counter, 1u);
    is_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {  // deterministic final pass: fixed order over the partials, in double
    __threadfence();
    double a = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) a += (double)((volatile float*)partial)[i];
    // reduce doubles via two float-free shuffles
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    __shared__ double shd[32];
    if ((threadIdx.x & 31) == 0) shd[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
      double tot = 0.0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += shd[w];
      *loss = (float)(tot / (double)n);
      *counter = 0;  // self-reset: workspace is reusable without a memset
    }
  }
}

__global__ void __launch_bounds__(256)
mse_bwd_kernel(const float* __restrict__ pred, const float* __restrict__ target,
               const float* __restrict__ grad_out, int64_t n, float* __restrict__ grad_pred, int vec) {
  const float g = __ldg(grad_out) * (2.0f / (float)n);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const int64_t nv = n >> 2;
    for (int64_t i = tid; i < nv; i += stride) {
      float4 p = ldg_stream_f4(pred + 4 * i), q = ldg_stream_f4(target + 4 * i), o;
      o.x = (p.x - q.x) * g; o.y = (p.y - q.y) * g; o.z = (p.z - q.z) * g; o.w = (p.w - q.w) * g;
      stg_stream_f4(grad_pred + 4 * i, o);
    }
    for (int64_t i = (nv << 2) + tid; i < n; i += stride) grad_pred[i] = (pred[i] - target[i]) * g;
  } else {
    for (int64_t i = tid; i < n; i += stride) grad_pred[i] = (pred[i] - target[i]) * g;
  }
}

static int mse_blocks(int64_t n) {
  int64_t b = (n / 4 + MSE_THREADS - 1) / MSE_THREADS;
  int64_t cap = (int64_t)num_sms() * 4;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_q_sample(const float* x0, const float* eps, const int64_t* t, const float* sqrt_acp,
                            const float* sqrt_1macp, int T, float* xt, int64_t B, int64_t per_sample,
                            int* err_flag, void* stream) {
  B2D_REQUIRE(x0 && eps && t && sqrt_acp && sqrt_1macp && xt, B2D_ERR_SHAPE, "q_sample: null pointer");
  B2D_REQUIRE(B >= 0 && per_sample >= 0 && T > 0, B2D_ERR_SHAPE, "q_sample: bad sizes B=%lld per_sample=%lld T=%d",
              (long long)B, (long long)per_sample, T);
  B2D_REQUIRE(B <= 65535, B2D_ERR_SHAPE, "q_sample: B=%lld exceeds 65535 samples per call", (long long)B);
  if (B == 0 || per_sample == 0) return B2D_OK;
  const bool vec = (per_sample % 4 == 0) && aligned16(x0) && aligned16(eps) && aligned16(xt);
  const int threads = 256;
  int64_t work = vec ? per_sample / 4 : per_sample;
  int64_t bx = (work + 2 * threads - 1) / (2 * threads);
  int64_t cap = ((int64_t)num_sms() * 8 + B - 1) / B;  // total CTAs ~ 8 per SM
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  dim3 grid((unsigned)bx, (unsigned)B);
  cudaStream_t s = (cudaStream_t)stream;
  if (vec) q_sample_kernel<true><<<grid, threads, 0, s>>>(x0, eps, t, sqrt_acp, sqrt_1macp, T, xt, per_sample, err_flag);
  else     q_sample_kernel<false><<<grid, threads, 0, s>>>(x0, eps, t, sqrt_acp, sqrt_1macp, T, xt, per_sample, err_flag);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_ddpm_step(const float* xt, const float* eps, const float* z, float s1m, float sqrt_acp,
                             float beta, float sqrt_alpha, float sigma, int use_noise, float* x_prev,
                             float* x0, int64_t n, void* stream) {
  StepCoef c{}; c.s1m = s1m; c.sqrt_acp = sqrt_acp; c.beta = beta; c.sqrt_alpha = sqrt_alpha;
  c.sigma = sigma; c.use_noise = use_noise;
  return launch_step<STEP_DDPM>(xt, eps, z, c, x_prev, x0, n, stream);
}
extern "C" int b2d_ddpm_v2_step(const float* xt, const float* eps, const float* z, float s1m,
                                float sqrt_alpha, float sqrt_acp, float sigma, float* x_prev, float* x0,
                                int64_t n, void* stream) {
  StepCoef c{}; c.s1m = s1m; c.sqrt_alpha = sqrt_alpha; c.sqrt_acp = sqrt_acp; c.sigma = sigma; c.use_noise = 1;
  return launch_step<STEP_V2>(xt, eps, z, c, x_prev, x0, n, stream);
}
extern "C" int b2d_cosine_step(const float* xt, const float* eps, const float* z, float s1m, float sqrt_acp,
                               float sigma, int use_noise, float* x_prev, float* mean_out, int64_t n,
                               void* stream) {
  StepCoef c{}; c.s1m = s1m; c.sqrt_acp = sqrt_acp; c.sigma = sigma; c.use_noise = use_noise;
  return launch_step<STEP_COS>(xt, eps, z, c, x_prev, mean_out, n, stream);
}
extern "C" int b2d_ddim_step(const float* xt, const float* eps, float s1m_t, float sqrt_acp_t,
                             float sqrt_acp_prev, float s1m_prev, float* x_prev, float* x0, int64_t n,
                             void* stream) {
  StepCoef c{}; c.s1m = s1m_t; c.sqrt_acp = sqrt_acp_t; c.sqrt_acp_prev = sqrt_acp_prev; c.s1m_prev = s1m_prev;
  return launch_step<STEP_DDIM>(xt, eps, nullptr, c, x_prev, x0, n, stream);
}

extern "C" size_t b2d_mse_workspace_bytes(int64_t n) {
  (void)n;
  return (size_t)(148 * 4 * 2) * sizeof(float) + 256;  // partials (generous) + counter
}
extern "C" int b2d_mse_fwd(const float* pred, const float* target, int64_t n, void* workspace,
                           size_t workspace_bytes, float* loss, void* stream) {
  B2D_REQUIRE(pred && target && loss && workspace, B2D_ERR_SHAPE, "mse_fwd: null pointer");
  B2D_REQUIRE(n > 0, B2D_ERR_SHAPE, "mse_fwd: n must be positive");
  int blocks = mse_blocks(n);
  B2D_REQUIRE(workspace_bytes >= 256 + (size_t)blocks * sizeof(float), B2D_ERR_WORKSPACE,
              "mse_fwd: workspace too small (%zu)", workspace_bytes);
  // layout: [0,256) counter (must be zero on first use; kernel self-resets), then partials
  unsigned int* counter = (unsigned int*)workspace;
  float* partial = (float*)((char*)workspace + 256);
  int vec = aligned16(pred) && aligned16(target);
  mse_fwd_kernel<<<blocks, MSE_THREADS, 0, (cudaStream_t)stream>>>(pred, target, n, partial, counter, loss, vec);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_mse_bwd(const float* pred, const float* target, const float* grad_out, int64_t n,
                           float* grad_pred, void* stream) {
  B2D_REQUIRE(pred && target && grad_out && grad_pred, B2D_ERR_SHAPE, "mse_bwd: null pointer");
  B2D_REQUIRE(n > 0, B2D_ERR_SHAPE, "mse_bwd: n must be positive");
  int vec = aligned16(pred) && aligned16(target) && aligned16(grad_pred);
  int64_t blocks = (n / 4 + 255) / 256;
  int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  mse_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(pred, target, grad_out, n, grad_pred, vec);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
