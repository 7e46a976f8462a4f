// optimizer.cu -- multi-tensor Adam: the optimiser step of train.py:102,134
// (`torch.optim.Adam(model.parameters(), 1e-2)` ... `optimizer.step()`) as ONE launch over every parameter tensor.
// HBM bound: 16 B read + 12 B written per parameter (p, g, m, v in; p, m, v out), fp32 throughout.
#include "common.cuh"

namespace b2d {

struct AdamScalars {
  float lr_over_bc1;   // lr / (1 - beta1^step)
  float inv_bc2_sqrt;  // 1 / sqrt(1 - beta2^step)
  float one_minus_beta1, beta2, one_minus_beta2, eps, weight_decay;   // 1 - beta rounded from double, as torch does
};

// The update in the order torch.optim.Adam's reference implementation (_single_tensor_adam) evaluates it:
//   g += wd * p;  m = lerp(m, g, 1 - b1);  v = v * b2 + (1 - b2) * g * g;
//   denom = sqrt(v) / sqrt(bc2) + eps;  p += -(lr / bc1) * (m / denom)
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamScalars& s) {
  if (s.weight_decay != 0.f) g = fmaf(s.weight_decay, p, g);
  m = fmaf(s.one_minus_beta1, g - m, m);
  v = fmaf(s.one_minus_beta2, g * g, v * s.beta2);
  const float denom = sqrtf(v) * s.inv_bc2_sqrt + s.eps;
  p = fmaf(-s.lr_over_bc1, __fdividef(m, denom), p);
}

constexpr int ADAM_THREADS = 512;

// one CTA per chunk (a run of <= chunk_elems consecutive elements of one tensor); grid = number of chunks
__global__ void __launch_bounds__(ADAM_THREADS)
adam_multi_tensor_kernel(const B2dAdamChunk* __restrict__ chunks, AdamScalars s, const int* __restrict__ found_overflow) {
  if (found_overflow != nullptr && *found_overflow != 0) return;      // skipped step (overflowed 16-bit gradients)
  const B2dAdamChunk c = chunks[blockIdx.x];
  float* __restrict__ p = c.param;
  const float* __restrict__ g = c.grad;
  float* __restrict__ m = c.exp_avg;
  float* __restrict__ v = c.exp_avg_sq;
  const int n = (int)c.n;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  int done = 0;
  if (vec) {
    const int n4 = n >> 2;
    // two independent 16-byte loads per array in flight per thread
    int i = threadIdx.x;
    for (; i + ADAM_THREADS < n4; i += 2 * ADAM_THREADS) {
      float4 p0 = reinterpret_cast<float4*>(p)[i], p1 = reinterpret_cast<float4*>(p)[i + ADAM_THREADS];
      const float4 g0 = ldg_stream_f4(g + 4 * i), g1 = ldg_stream_f4(g + 4 * (i + ADAM_THREADS));
      float4 m0 = reinterpret_cast<float4*>(m)[i], m1 = reinterpret_cast<float4*>(m)[i + ADAM_THREADS];
      float4 v0 = reinterpret_cast<float4*>(v)[i], v1 = reinterpret_cast<float4*>(v)[i + ADAM_THREADS];
      adam_one(p0.x, g0.x, m0.x, v0.x, s); adam_one(p0.y, g0.y, m0.y, v0.y, s);
      adam_one(p0.z, g0.z, m0.z, v0.z, s); adam_one(p0.w, g0.w, m0.w, v0.w, s);
      adam_one(p1.x, g1.x, m1.x, v1.x, s); adam_one(p1.y, g1.y, m1.y, v1.y, s);
      adam_one(p1.z, g1.z, m1.z, v1.z, s); adam_one(p1.w, g1.w, m1.w, v1.w, s);
      reinterpret_cast<float4*>(p)[i] = p0; reinterpret_cast<float4*>(p)[i + ADAM_THREADS] = p1;
      reinterpret_cast<float4*>(m)[i] = m0; reinterpret_cast<float4*>(m)[i + ADAM_THREADS] = m1;
      reinterpret_cast<float4*>(v)[i] = v0; reinterpret_cast<float4*>(v)[i + ADAM_THREADS] = v1;
    }
    for (; i < n4; i += ADAM_THREADS) {
      float4 p0 = reinterpret_cast<float4*>(p)[i];
      const float4 g0 = ldg_stream_f4(g + 4 * i);
      float4 m0 = reinterpret_cast<float4*>(m)[i];
      float4 v0 = reinterpret_cast<float4*>(v)[i];
      adam_one(p0.x, g0.x, m0.x, v0.x, s); adam_one(p0.y, g0.y, m0.y, v0.y, s);
      adam_one(p0.z, g0.z, m0.z, v0.z, s); adam_one(p0.w, g0.w, m0.w, v0.w, s);
      reinterpret_cast<float4*>(p)[i] = p0;
      reinterpret_cast<float4*>(m)[i] = m0;
      reinterpret_cast<float4*>(v)[i] = v0;
    }
    done = n4 << 2;
  }
  for (int i = done + threadIdx.x; i < n; i += ADAM_THREADS) {
    float pp = p[i], mm = m[i], vv = v[i];
    adam_one(pp, g[i], mm, vv, s);
    p[i] = pp; m[i] = mm; v[i] = vv;
  }
}

}  // namespace b2d

extern "C" int b2d_adam_step(const B2dAdamChunk* chunks_dev, int n_chunks, double lr, double beta1_d, double beta2_d, float eps,
                             float weight_decay, double bias_correction1, double bias_correction2, void* stream) {
  return b2d_adam_step_guarded(chunks_dev, n_chunks, lr, beta1_d, beta2_d, eps, weight_decay, bias_correction1,
                               bias_correction2, nullptr, stream);
}

extern "C" int b2d_adam_step_guarded(const B2dAdamChunk* chunks_dev, int n_chunks, double lr, double beta1_d, double beta2_d,
                                     float eps, float weight_decay, double bias_correction1, double bias_correction2,
                                     const int32_t* found_overflow, void* stream) {
  using namespace b2d;
  if (n_chunks == 0) return B2D_OK;
  B2D_REQUIRE(chunks_dev != nullptr && n_chunks > 0, B2D_ERR_SHAPE, "adam_step: empty chunk table");
  B2D_REQUIRE(bias_correction1 > 0.0 && bias_correction2 > 0.0, B2D_ERR_SHAPE, "adam_step: bias corrections must be > 0");
  B2D_REQUIRE(aligned16(chunks_dev), B2D_ERR_ALIGN, "adam_step: chunk table must be 16-byte aligned");
  AdamScalars s;
  s.lr_over_bc1 = (float)(lr / bias_correction1);
  s.inv_bc2_sqrt = (float)(1.0 / sqrt(bias_correction2));
  s.one_minus_beta1 = (float)(1.0 - (double)beta1_d); s.beta2 = (float)beta2_d; s.one_minus_beta2 = (float)(1.0 - beta2_d);
  s.eps = eps; s.weight_decay = weight_decay;
  adam_multi_tensor_kernel<<<n_chunks, ADAM_THREADS, 0, (cudaStream_t)stream>>>(chunks_dev, s, found_overflow);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
