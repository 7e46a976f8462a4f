// groupnorm.cu -- GroupNorm32(32, C) (+ SiLU, + dropout) forward and backward on pixel-major activations
// (forward tensors x / y are fp16, gradient tensors dy / dx / add_* are bf16; pointers are typed
// __half* only as a 16-bit placeholder) (utils.py:54-56,130-137; used at unet.py:194-198,218-225,297,624-628).
// Statistics are accumulated in fp32 per thread and fp64 across CTAs, as the reference computes
// GroupNorm in fp32 regardless of activation dtype.  The input may be the channel concatenation
// of two tensors (decoder skip connections, unet.py:672) -- it is never materialised.
//
// HBM-bound kernels: every thread moves 16-byte vectors (8 bf16 channels of one pixel).
#include "common.cuh"
#include <stdlib.h>

namespace b2d {

struct GnSrc {
  const __half* x1; long long ld1; int C1;
  const __half* x2; long long ld2; int C2;
};

__device__ __forceinline__ const __half* src_ptr(const GnSrc& s, long long pix, int c) {
  return (c < s.C1) ? s.x1 + pix * s.ld1 + c : s.x2 + pix * s.ld2 + (c - s.C1);
}

__device__ __forceinline__ void load8(const __half* p, float* v) {
  uint4 u = ldg_nc_u4(p);
  float2 a = unpack_f16x2(u.x), b = unpack_f16x2(u.y), c = unpack_f16x2(u.z), d = unpack_f16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void store8(__half* p, const float* v) {
  uint4 u;
  u.x = pack_f16x2(v[0], v[1]); u.y = pack_f16x2(v[2], v[3]);
  u.z = pack_f16x2(v[4], v[5]); u.w = pack_f16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// HBM-bound kernels below keep several independent 16-byte loads in flight per thread (UNROLL pixels per
// iteration): with one load per thread the first version sustained ~2 TB/s (16 KB in flight per SM).

// keep-mask for 8 consecutive elements starting at element index e (multiple of 8)
__device__ __forceinline__ void dropout_keep8(unsigned long long seed, unsigned long long e, float p, bool* keep) {
  const unsigned long long blk = e >> 3;
  uint4 a = philox4((uint32_t)seed, (uint32_t)(seed >> 32), (uint32_t)blk, (uint32_t)(blk >> 32));
  const uint32_t w[4] = {a.x, a.y, a.z, a.w};
  const uint32_t thr = (uint32_t)(p * 65536.0f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    keep[2 * i] = (w[i] & 0xFFFFu) >= thr;
    keep[2 * i + 1] = (w[i] >> 16) >= thr;
  }
}

// ------------------------------------------------------------------ forward: statistics
// grid (chunks, N); block = V * rows threads (V = Ctot/8 vectors per pixel).
// sums: double [N][32][2] (sum, sum of squares), zeroed by the caller.
__global__ void __launch_bounds__(256)
gn_stats_kernel(GnSrc s, int HW, int Ctot, double* __restrict__ sums) {
  extern __shared__ float sh[];  // [2*Ctot]
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V;
  const int prow = threadIdx.x / V;
  const int rows = blockDim.x / V;
  const int n = blockIdx.y;
  for (int i = threadIdx.x; i < 2 * Ctot; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  float a[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = 0.f; q[i] = 0.f; }
  const int c = v * 8;
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 4 * step) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix0 + u * step < HW) raw[u] = ldg_nc_u4(src_ptr(s, (long long)n * HW + pix0 + u * step, c));
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (pix0 + u * step < HW) {
        float x[8];
        unpack8(raw[u], x);
#pragma unroll
        for (int i = 0; i < 8; ++i) { a[i] += x[i]; q[i] = fmaf(x[i], x[i], q[i]); }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { atomicAdd(&sh[c + i], a[i]); atomicAdd(&sh[Ctot + c + i], q[i]); }
  __syncthreads();
  const int cpg = Ctot >> 5;
  if (threadIdx.x < 32) {
    float sa = 0.f, sq = 0.f;
    for (int i = 0; i < cpg; ++i) { sa += sh[threadIdx.x * cpg + i]; sq += sh[Ctot + threadIdx.x * cpg + i]; }
    atomicAdd(&sums[((long long)n * 32 + threadIdx.x) * 2 + 0], (double)sa);
    atomicAdd(&sums[((long long)n * 32 + threadIdx.x) * 2 + 1], (double)sq);
  }
}

// ------------------------------------------------------------------ forward: normalise (+SiLU, +dropout)
// act: 0 = none, 1 = SiLU.  Writes y (N,HW,Ctot) contiguous and mean/rstd [N][32] for the backward.
__global__ void __launch_bounds__(256)
gn_apply_kernel(GnSrc s, int HW, int Ctot, const double* __restrict__ sums, const float* __restrict__ gamma,
                const float* __restrict__ beta, float eps, int act, float drop_p, unsigned long long seed,
                __half* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  __shared__ float s_mean[32], s_rstd[32];
  const int n = blockIdx.y;
  const int cpg = Ctot >> 5;
  if (threadIdx.x < 32) {
    const double cnt = (double)cpg * (double)HW;
    const double m = sums[((long long)n * 32 + threadIdx.x) * 2] / cnt;
    double var = sums[((long long)n * 32 + threadIdx.x) * 2 + 1] / cnt - m * m;
    if (var < 0.0) var = 0.0;
    const float r = (float)(1.0 / sqrt(var + (double)eps));
    s_mean[threadIdx.x] = (float)m;
    s_rstd[threadIdx.x] = r;
    if (blockIdx.x == 0) { mean_out[n * 32 + threadIdx.x] = (float)m; rstd_out[n * 32 + threadIdx.x] = r; }
  }
  __syncthreads();
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V;
  const int prow = threadIdx.x / V;
  const int rows = blockDim.x / V;
  const int c = v * 8;
  float g[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float gm = __ldg(gamma + c + i), rs_ = s_rstd[(c + i) / cpg];
    g[i] = gm * rs_;                                   // (x - mu) * rs * gamma + beta = x * g + b
    b[i] = fmaf(-s_mean[(c + i) / cpg], g[i], __ldg(beta + c + i));
  }
  const float keep_scale = drop_p > 0.f ? 1.0f / (1.0f - drop_p) : 1.0f;
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 4 * step) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix0 + u * step < HW) raw[u] = ldg_nc_u4(src_ptr(s, (long long)n * HW + pix0 + u * step, c));
#pragma unroll
    for (int u = 0; u < 4; ++u) {
    const int pix = pix0 + u * step;
    if (pix >= HW) continue;
    const long long gp = (long long)n * HW + pix;
    float x[8];
    unpack8(raw[u], x);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float z = fmaf(x[i], g[i], b[i]);               // g, b hold the folded per-channel scale / shift
      x[i] = act ? silu_f(z) : z;
    }
    if (drop_p > 0.f) {
      bool keep[8];
      dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = keep[i] ? x[i] * keep_scale : 0.f;
    }
    store8_fmt(y + gp * Ctot + c, x, FMT_F16);
    }
  }
}

// ------------------------------------------------------------------ backward: per-(sample, channel) sums
// ab: fp32 [N][Ctot][2] = (sum dyh, sum dyh * xhat) with dyh = dy * act'(z) (* dropout mask), zeroed by caller.
__global__ void __launch_bounds__(256, 3)
gn_bwd_reduce_kernel(GnSrc s, const __half* __restrict__ dy, int HW, int Ctot,
                     const float* __restrict__ mean, const float* __restrict__ rstd,
                     const float* __restrict__ gamma, const float* __restrict__ beta, int act, float drop_p,
                     unsigned long long seed, float* __restrict__ ab) {
  extern __shared__ float sh[];  // [2*Ctot]
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V;
  const int prow = threadIdx.x / V;
  const int rows = blockDim.x / V;
  const int n = blockIdx.y;
  const int cpg = Ctot >> 5;
  for (int i = threadIdx.x; i < 2 * Ctot; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int c = v * 8;
  // z = xhat * gamma + beta = x * rg + bb;  sum dyh * xhat = rstd * (sum dyh * x) - mean * rstd * (sum dyh):
  // two folded constants per channel in the loop instead of four keeps the kernel at 3 blocks / SM
  float rg[8], bb[8], A[8], B[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float rs_ = __ldg(rstd + n * 32 + (c + i) / cpg), mu_ = __ldg(mean + n * 32 + (c + i) / cpg);
    rg[i] = rs_ * __ldg(gamma + c + i);
    bb[i] = fmaf(-mu_, rg[i], __ldg(beta + c + i));
    A[i] = 0.f; B[i] = 0.f;
  }
  const float keep_scale = drop_p > 0.f ? 1.0f / (1.0f - drop_p) : 1.0f;
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 2 * step) {
    uint4 rx[2], rd[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pix0 + u * step < HW) {
        const long long gp = (long long)n * HW + pix0 + u * step;
        rx[u] = ldg_nc_u4(src_ptr(s, gp, c));
        rd[u] = ldg_nc_u4(dy + gp * Ctot + c);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
    const int pix = pix0 + u * step;
    if (pix >= HW) continue;
    const long long gp = (long long)n * HW + pix;
    float x[8], d[8];
    unpack8(rx[u], x);
    unpack8(rd[u], d);
    if (drop_p > 0.f) {
      bool keep[8];
      dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
#pragma unroll
      for (int i = 0; i < 8; ++i) d[i] = keep[i] ? d[i] * keep_scale : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float dh = d[i];
      if (act) {
        const float z = fmaf(x[i], rg[i], bb[i]);
        const float sg = fast_sigmoid(z);
        dh *= sg * (1.0f + z * (1.0f - sg));
      }
      A[i] += dh; B[i] = fmaf(dh, x[i], B[i]);
    }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float rs_ = __ldg(rstd + n * 32 + (c + i) / cpg), mu_ = __ldg(mean + n * 32 + (c + i) / cpg);
    atomicAdd(&sh[c + i], A[i]);
    atomicAdd(&sh[Ctot + c + i], rs_ * (B[i] - mu_ * A[i]));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < Ctot; i += blockDim.x) {
    atomicAdd(&ab[((long long)n * Ctot + i) * 2 + 0], sh[i]);
    atomicAdd(&ab[((long long)n * Ctot + i) * 2 + 1], sh[Ctot + i]);
  }
}

// ------------------------------------------------------------------ backward: input gradient
// dx = rstd * (dyh*gamma - s1/m - xhat*s2/m) (+ add_a + add_b on source 1 channels, + nothing on source 2)
struct GnBwdOut {
  __half* dx1; long long ldd1;
  __half* dx2; long long ldd2;
  const __half* add_a; long long lda;
  const __half* add_b; long long ldb;
  int add_a_full;  // add_a spans all C1+C2 channels (also added to dx2)
  float* colsum; long long ldcs;  // optional: colsum[n][c] += sum over pixels of dx1[n, :, c] (c < C1), fp32
  int* overflow;                  // optional: += 1 per thread that saw a saturated (|v| >= 65504) or NaN gradient element
};

// Per-sample column sums of the dx1 values a thread produced (cs[8], channels c..c+7 of source 1), folded into
// o.colsum through a shared-memory stage: the consumer of dx1 as a convolution's output gradient needs exactly these
// sums for the bias (and timestep-embedding) gradients, and a separate pass would read the tensor from HBM again.
__device__ __forceinline__ void gn_colsum_flush(const GnBwdOut& o, float* shc, const float* cs, int C1, int c, int n,
                                                bool first) {
  for (int i = threadIdx.x; i < C1; i += blockDim.x) shc[i] = 0.f;
  __syncthreads();
  if (first) {
#pragma unroll
    for (int i = 0; i < 8; ++i) atomicAdd(&shc[c + i], cs[i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C1; i += blockDim.x) atomicAdd(&o.colsum[(long long)n * o.ldcs + i], shc[i]);
}

template <bool CS>   // CS: also accumulate the per-sample column sums of dx1 (8 more live registers: 2 CTAs / SM)
__global__ void __launch_bounds__(256, CS ? 2 : 3)
gn_bwd_apply_kernel(GnSrc s, const __half* __restrict__ dy, int HW, int Ctot,
                    const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ gamma, const float* __restrict__ beta, int act, float drop_p,
                    unsigned long long seed, const float* __restrict__ ab, GnBwdOut o) {
  __shared__ float s1[32], s2[32];
  extern __shared__ float shc[];   // [C1] when o.colsum
  const int n = blockIdx.y;
  const int cpg = Ctot >> 5;
  if (threadIdx.x < 32) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < cpg; ++i) {
      const int ch = threadIdx.x * cpg + i;
      const float gm = __ldg(gamma + ch);
      a = fmaf(gm, ab[((long long)n * Ctot + ch) * 2 + 0], a);
      b = fmaf(gm, ab[((long long)n * Ctot + ch) * 2 + 1], b);
    }
    const float inv_m = 1.0f / ((float)cpg * (float)HW);
    s1[threadIdx.x] = a * inv_m;
    s2[threadIdx.x] = b * inv_m;
  }
  __syncthreads();
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V;
  const int prow = threadIdx.x / V;
  const int rows = blockDim.x / V;
  const int c = v * 8;
  // z = x * rg + bb;  dx = rstd * (dyh * gamma - s1 - xhat * s2) = dyh * rg + x * pp + qq
  float rg[8], bb[8], pp[8], qq[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int gi = (c + i) / cpg;
    const float rs_ = __ldg(rstd + n * 32 + gi), mu_ = __ldg(mean + n * 32 + gi);
    rg[i] = rs_ * __ldg(gamma + c + i);
    bb[i] = fmaf(-mu_, rg[i], __ldg(beta + c + i));
    pp[i] = -rs_ * rs_ * s2[gi];
    qq[i] = -mu_ * pp[i] - rs_ * s1[gi];
  }
  const float keep_scale = drop_p > 0.f ? 1.0f / (1.0f - drop_p) : 1.0f;
  const bool first = c < s.C1;
  const bool has_a = o.add_a && (first || o.add_a_full);
  const bool has_b = o.add_b && first;
  float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float amax = 0.f;
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 2 * step) {
    uint4 rx[2], rd[2], ra[2], rb[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pix0 + u * step < HW) {
        const long long gp = (long long)n * HW + pix0 + u * step;
        rx[u] = ldg_nc_u4(src_ptr(s, gp, c));
        rd[u] = ldg_nc_u4(dy + gp * Ctot + c);
        if (has_a) ra[u] = ldg_nc_u4(o.add_a + gp * o.lda + c);
        if (has_b) rb[u] = ldg_nc_u4(o.add_b + gp * o.ldb + c);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
    const int pix = pix0 + u * step;
    if (pix >= HW) continue;
    const long long gp = (long long)n * HW + pix;
    float x[8], d[8];
    unpack8(rx[u], x);
    unpack8(rd[u], d);
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(d[i]));
    if (drop_p > 0.f) {
      bool keep[8];
      dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
#pragma unroll
      for (int i = 0; i < 8; ++i) d[i] = keep[i] ? d[i] * keep_scale : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float dh = d[i];
      if (act) {
        const float z = fmaf(x[i], rg[i], bb[i]);
        const float sg = fast_sigmoid(z);
        dh *= sg * (1.0f + z * (1.0f - sg));
      }
      x[i] = fmaf(dh, rg[i], fmaf(x[i], pp[i], qq[i]));
    }
    if (has_a) { float t[8]; unpack8(ra[u], t);
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] += t[i]; }
    if (has_b) { float t[8]; unpack8(rb[u], t);
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] += t[i]; }
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(x[i]));
    if (first) {
      store8(o.dx1 + gp * o.ldd1 + c, x);
      if (CS) {
#pragma unroll
        for (int i = 0; i < 8; ++i) cs[i] += x[i];
      }
    } else {
      store8(o.dx2 + gp * o.ldd2 + (c - s.C1), x);
    }
    }
  }
  if (o.overflow != nullptr && !(amax < 65504.f)) atomicAdd(o.overflow, 1);   // saturated (or NaN) 16-bit gradient seen
  if (CS) gn_colsum_flush(o, shc, cs, s.C1, c, n, first);
}

// dgamma[c] = sum_n ab[n][c][1], dbeta[c] = sum_n ab[n][c][0]  (accumulated into fp32 grads)
// block = 32 channels x 8 sample lanes (the one-thread-per-channel version walked N samples serially: ~30 us)
__global__ void __launch_bounds__(256)
gn_param_grad_kernel(const float* __restrict__ ab, int N, int Ctot, float* __restrict__ dgamma,
                     float* __restrict__ dbeta) {
  __shared__ float sa[8][33], sb[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float a = 0.f, b = 0.f;
  if (c < Ctot) {
    for (int n0 = ty; n0 < N; n0 += 32) {
      float2 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        t[u] = (n0 + 8 * u < N) ? __ldg(reinterpret_cast<const float2*>(ab + ((long long)(n0 + 8 * u) * Ctot + c) * 2))
                                : make_float2(0.f, 0.f);
      a += (t[0].x + t[1].x) + (t[2].x + t[3].x);
      b += (t[0].y + t[1].y) + (t[2].y + t[3].y);
    }
  }
  sa[ty][tx] = a; sb[ty][tx] = b;
  __syncthreads();
  if (ty == 0 && c < Ctot) {
    float ta = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { ta += sa[i][tx]; tb += sb[i][tx]; }
    dbeta[c] += ta;
    dgamma[c] += tb;
  }
}

// ====================================================================== single-pass (cluster) variants
// The two-kernel forms above read x (and dy) from HBM twice: statistics, then apply.  Here one thread-block CLUSTER
// owns one sample: each CTA stages its pixel range of x in shared memory with cp.async (every thread re-reads only the
// 16-byte vectors it loaded itself, so no CTA barrier guards the tile), the per-group partial sums are exchanged through
// distributed shared memory (64 floats per CTA) across one cluster barrier, and the apply pass reads x from shared
// memory.  HBM traffic: forward 1 read + 1 write (was 2 + 1); backward x once, dy once from HBM and once more from L2
// (it was read microseconds earlier by the same SM; the in-flight footprint is < 30 MB), dx once (was 2 + 2 + 1).
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ float ld_dsmem_f32(const float* local_ptr, uint32_t rank) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(local_ptr);
  uint32_t ra;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
  return v;
}
__device__ __forceinline__ void cp_async16(void* smem, const void* g) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(a), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// grid (CL, N), cluster (CL, 1, 1); block = V * rows threads; dynamic smem = ppc * Ctot * 2 + (2 * Ctot + 64) * 4 bytes
__global__ void __launch_bounds__(256)
gn_fwd_fused_kernel(GnSrc s, int HW, int Ctot, int ppc, const float* __restrict__ gamma, const float* __restrict__ beta,
                    float eps, int act, float drop_p, unsigned long long seed, __half* __restrict__ y,
                    float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  extern __shared__ __align__(16) unsigned char gn_smem[];
  uint4* tile = reinterpret_cast<uint4*>(gn_smem);                                   // [ppc][V]
  float* sh = reinterpret_cast<float*>(gn_smem + (size_t)ppc * Ctot * 2);            // [2 * Ctot]
  float* part = sh + 2 * Ctot;                                                       // [32][2]
  __shared__ float s_mean[32], s_rstd[32];
  const int CL = gridDim.x, rank = blockIdx.x, n = blockIdx.y;
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V, prow = threadIdx.x / V, rows = blockDim.x / V;
  const int c = v * 8;
  const int cpg = Ctot >> 5;
  const int p0 = rank * ppc;
  const int p1 = (p0 + ppc < HW) ? p0 + ppc : HW;
  for (int pix = p0 + prow; pix < p1; pix += rows)
    cp_async16(&tile[(pix - p0) * V + v], src_ptr(s, (long long)n * HW + pix, c));
  for (int i = threadIdx.x; i < 2 * Ctot; i += blockDim.x) sh[i] = 0.f;
  cp_async_wait_all();
  __syncthreads();
  float a[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = 0.f; q[i] = 0.f; }
  for (int pix = p0 + prow; pix < p1; pix += rows) {
    float x[8];
    unpack8(tile[(pix - p0) * V + v], x);
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] += x[i]; q[i] = fmaf(x[i], x[i], q[i]); }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { atomicAdd(&sh[c + i], a[i]); atomicAdd(&sh[Ctot + c + i], q[i]); }
  __syncthreads();
  if (threadIdx.x < 32) {
    float sa = 0.f, sq = 0.f;
    for (int i = 0; i < cpg; ++i) { sa += sh[threadIdx.x * cpg + i]; sq += sh[Ctot + threadIdx.x * cpg + i]; }
    part[threadIdx.x * 2] = sa;
    part[threadIdx.x * 2 + 1] = sq;
  }
  cluster_arrive();
  cluster_wait();                                       // every CTA's partial sums are visible cluster-wide
  if (threadIdx.x < 32) {
    double S = 0.0, Q = 0.0;
    for (int r = 0; r < CL; ++r) {
      S += (double)ld_dsmem_f32(&part[threadIdx.x * 2], r);
      Q += (double)ld_dsmem_f32(&part[threadIdx.x * 2 + 1], r);
    }
    const double cnt = (double)cpg * (double)HW;
    const double m = S / cnt;
    double var = Q / cnt - m * m;
    if (var < 0.0) var = 0.0;
    const float rs = (float)(1.0 / sqrt(var + (double)eps));
    s_mean[threadIdx.x] = (float)m;
    s_rstd[threadIdx.x] = rs;
    if (rank == 0) { mean_out[n * 32 + threadIdx.x] = (float)m; rstd_out[n * 32 + threadIdx.x] = rs; }
  }
  cluster_arrive();                                     // done with the peers' shared memory (waited on before exit)
  __syncthreads();
  float g[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float gm = __ldg(gamma + c + i), rs_ = s_rstd[(c + i) / cpg];
    g[i] = gm * rs_;
    b[i] = fmaf(-s_mean[(c + i) / cpg], g[i], __ldg(beta + c + i));
  }
  const float keep_scale = drop_p > 0.f ? 1.0f / (1.0f - drop_p) : 1.0f;
  for (int pix = p0 + prow; pix < p1; pix += rows) {
    const long long gp = (long long)n * HW + pix;
    float x[8];
    unpack8(tile[(pix - p0) * V + v], x);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float z = fmaf(x[i], g[i], b[i]);
      x[i] = act ? silu_f(z) : z;
    }
    if (drop_p > 0.f) {
      bool keep[8];
      dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
#pragma unroll
      for (int i = 0; i < 8; ++i) x[i] = keep[i] ? x[i] * keep_scale : 0.f;
    }
    store8_fmt(y + gp * Ctot + c, x, FMT_F16);
  }
  cluster_wait();
}

template <bool CS>
__global__ void __launch_bounds__(256, CS ? 2 : 3)
gn_bwd_fused_kernel(GnSrc s, const __half* __restrict__ dy, int HW, int Ctot, int ppc,
                    const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ gamma, const float* __restrict__ beta, int act, float drop_p,
                    unsigned long long seed, float* __restrict__ ab, GnBwdOut o) {
  extern __shared__ __align__(16) unsigned char gn_smem[];
  uint4* tile = reinterpret_cast<uint4*>(gn_smem);                                   // x, [ppc][V]
  float* sh = reinterpret_cast<float*>(gn_smem + (size_t)ppc * Ctot * 2);            // [2 * Ctot]
  float* part = sh + 2 * Ctot;                                                       // [32][2]
  __shared__ float s1[32], s2[32];
  const int CL = gridDim.x, rank = blockIdx.x, n = blockIdx.y;
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V, prow = threadIdx.x / V, rows = blockDim.x / V;
  const int c = v * 8;
  const int cpg = Ctot >> 5;
  const int p0 = rank * ppc;
  const int p1 = (p0 + ppc < HW) ? p0 + ppc : HW;
  for (int pix = p0 + prow; pix < p1; pix += rows)
    cp_async16(&tile[(pix - p0) * V + v], src_ptr(s, (long long)n * HW + pix, c));
  for (int i = threadIdx.x; i < 2 * Ctot; i += blockDim.x) sh[i] = 0.f;
  float rg[8], bb[8], pp[8], qq[8];                     // pp / qq hold the running sums A / B in the first pass
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int gi = (c + i) / cpg;
    const float rs_ = __ldg(rstd + n * 32 + gi), mu_ = __ldg(mean + n * 32 + gi);
    rg[i] = rs_ * __ldg(gamma + c + i);
    bb[i] = fmaf(-mu_, rg[i], __ldg(beta + c + i));
    pp[i] = 0.f; qq[i] = 0.f;
  }
  const float keep_scale = drop_p > 0.f ? 1.0f / (1.0f - drop_p) : 1.0f;
  cp_async_wait_all();
  __syncthreads();
  for (int pix0 = p0 + prow; pix0 < p1; pix0 += 4 * rows) {
    uint4 rd[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix0 + u * rows < p1) rd[u] = ldg_nc_u4(dy + ((long long)n * HW + pix0 + u * rows) * Ctot + c);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pix = pix0 + u * rows;
      if (pix >= p1) continue;
      const long long gp = (long long)n * HW + pix;
      float x[8], d[8];
      unpack8(tile[(pix - p0) * V + v], x);
      unpack8(rd[u], d);
      if (drop_p > 0.f) {
        bool keep[8];
        dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] = keep[i] ? d[i] * keep_scale : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float dh = d[i];
        if (act) {
          const float z = fmaf(x[i], rg[i], bb[i]);
          const float sg = fast_sigmoid(z);
          dh *= sg * (1.0f + z * (1.0f - sg));
        }
        pp[i] += dh; qq[i] = fmaf(dh, x[i], qq[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float rs_ = __ldg(rstd + n * 32 + (c + i) / cpg), mu_ = __ldg(mean + n * 32 + (c + i) / cpg);
    atomicAdd(&sh[c + i], pp[i]);
    atomicAdd(&sh[Ctot + c + i], rs_ * (qq[i] - mu_ * pp[i]));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < Ctot; i += blockDim.x) {        // per-channel sums for dgamma / dbeta
    atomicAdd(&ab[((long long)n * Ctot + i) * 2 + 0], sh[i]);
    atomicAdd(&ab[((long long)n * Ctot + i) * 2 + 1], sh[Ctot + i]);
  }
  if (threadIdx.x < 32) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < cpg; ++i) {
      const int ch = threadIdx.x * cpg + i;
      const float gm = __ldg(gamma + ch);
      a = fmaf(gm, sh[ch], a);
      b = fmaf(gm, sh[Ctot + ch], b);
    }
    part[threadIdx.x * 2] = a;
    part[threadIdx.x * 2 + 1] = b;
  }
  cluster_arrive();
  cluster_wait();
  if (threadIdx.x < 32) {
    float a = 0.f, b = 0.f;
    for (int r = 0; r < CL; ++r) {
      a += ld_dsmem_f32(&part[threadIdx.x * 2], r);
      b += ld_dsmem_f32(&part[threadIdx.x * 2 + 1], r);
    }
    const float inv_m = 1.0f / ((float)cpg * (float)HW);
    s1[threadIdx.x] = a * inv_m;
    s2[threadIdx.x] = b * inv_m;
  }
  cluster_arrive();
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int gi = (c + i) / cpg;
    const float rs_ = __ldg(rstd + n * 32 + gi), mu_ = __ldg(mean + n * 32 + gi);
    pp[i] = -rs_ * rs_ * s2[gi];
    qq[i] = -mu_ * pp[i] - rs_ * s1[gi];
  }
  const bool first = c < s.C1;
  const bool has_a = o.add_a && (first || o.add_a_full);
  const bool has_b = o.add_b && first;
  float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float amax = 0.f;
  for (int pix0 = p0 + prow; pix0 < p1; pix0 += 2 * rows) {
    uint4 rd[2], ra[2], rb[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pix0 + u * rows < p1) {
        const long long gp = (long long)n * HW + pix0 + u * rows;
        rd[u] = ldg_nc_u4(dy + gp * Ctot + c);
        if (has_a) ra[u] = ldg_nc_u4(o.add_a + gp * o.lda + c);
        if (has_b) rb[u] = ldg_nc_u4(o.add_b + gp * o.ldb + c);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int pix = pix0 + u * rows;
      if (pix >= p1) continue;
      const long long gp = (long long)n * HW + pix;
      float x[8], d[8];
      unpack8(tile[(pix - p0) * V + v], x);
      unpack8(rd[u], d);
      if (drop_p > 0.f) {
        bool keep[8];
        dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i] = keep[i] ? d[i] * keep_scale : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float dh = d[i];
        if (act) {
          const float z = fmaf(x[i], rg[i], bb[i]);
          const float sg = fast_sigmoid(z);
          dh *= sg * (1.0f + z * (1.0f - sg));
        }
        x[i] = fmaf(dh, rg[i], fmaf(x[i], pp[i], qq[i]));
      }
      if (has_a) { float t[8]; unpack8(ra[u], t);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] += t[i]; }
      if (has_b) { float t[8]; unpack8(rb[u], t);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] += t[i]; }
#pragma unroll
      for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fmaxf(fabsf(x[i]), fabsf(d[i])));
      if (first) {
        store8(o.dx1 + gp * o.ldd1 + c, x);
        if (CS) {
#pragma unroll
          for (int i = 0; i < 8; ++i) cs[i] += x[i];
        }
      } else {
        store8(o.dx2 + gp * o.ldd2 + (c - s.C1), x);
      }
    }
  }
  if (o.overflow != nullptr && !(amax < 65504.f)) atomicAdd(o.overflow, 1);
  cluster_wait();                                       // after it no peer reads this CTA's `part`; `sh` is free for reuse
  if (CS) gn_colsum_flush(o, sh, cs, s.C1, c, n, first);
}

// ====================================================================== producer-fused statistics (engine path)
// The convolution that produces a tensor emits, from its fp32 epilogue values, the per-(sample, channel pair) sum and
// sum of squares (conv_gemm.cu EPI_STATS); the GroupNorm that consumes the tensor -- possibly as one half of a skip
// concatenation -- forms its 32 group moments from those pairs: forward = ONE pass (read x, write y).  The
// data-gradient convolution that produces dL/dy emits dyh = dy * act'(z) * mask and (sum dyh, sum dyh * x) per
// (sample, channel) (EPI_GNB); backward = ONE pass (read x, dyh [, add_a, add_b], write dx) + the parameter sums.

// st[n][c/2][0..1] += (sum, sum of squares) of channels (c, c+1) over the pixels of sample n.  grid (chunks, N).
__global__ void __launch_bounds__(256)
pair_stats_kernel(const __half* __restrict__ x, long long ld, int C, int HW, double* __restrict__ st) {
  extern __shared__ float sh[];  // [C]: pair p -> sh[2p], sh[2p + 1]
  const int V = C >> 3;
  const int v = threadIdx.x % V, prow = threadIdx.x / V, rows = blockDim.x / V;
  const int n = blockIdx.y;
  for (int i = threadIdx.x; i < C; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  float a[4] = {0.f, 0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f};
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 4 * step) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix0 + u * step < HW) raw[u] = ldg_nc_u4(x + ((long long)n * HW + pix0 + u * step) * ld + v * 8);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (pix0 + u * step < HW) {
        float f[8];
        unpack8(raw[u], f);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          a[i] += f[2 * i] + f[2 * i + 1];
          q[i] = fmaf(f[2 * i], f[2 * i], fmaf(f[2 * i + 1], f[2 * i + 1], q[i]));
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) { atomicAdd(&sh[v * 8 + 2 * i], a[i]); atomicAdd(&sh[v * 8 + 2 * i + 1], q[i]); }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) atomicAdd(&st[(long long)n * C + i], (double)sh[i]);
}

// group moments of sample n from the pair sums of the (up to two) source tensors
__device__ __forceinline__ void gn_group_moments(const double* __restrict__ st1, const double* __restrict__ st2, int C1,
                                                 int C2, int rps, int n, int g, int cpg, int HW, float eps, float* mean,
                                                 float* rstd) {
  double S = 0.0, Q = 0.0;
  const int p0 = g * (cpg >> 1), p1 = p0 + (cpg >> 1), h1 = C1 >> 1, h2 = C2 >> 1;
  for (int r = 0; r < rps; ++r) {
    const long long row = (long long)n * rps + r;
    for (int pi = p0; pi < p1; ++pi) {
      const double* s = pi < h1 ? st1 + (row * h1 + pi) * 2 : st2 + (row * h2 + (pi - h1)) * 2;
      S += s[0]; Q += s[1];
    }
  }
  const double cnt = (double)cpg * (double)HW;
  const double m = S / cnt;
  double var = Q / cnt - m * m;
  if (var < 0.0) var = 0.0;
  *mean = (float)m;
  *rstd = (float)(1.0 / sqrt(var + (double)eps));
}

__global__ void __launch_bounds__(256)
gn_apply2_kernel(GnSrc s, int HW, int Ctot, const double* __restrict__ st1, const double* __restrict__ st2, int rps,
                 const float* __restrict__ gamma, const float* __restrict__ beta, float eps, int act, float drop_p,
                 unsigned long long seed, __half* __restrict__ y, float* __restrict__ mean_out,
                 float* __restrict__ rstd_out, unsigned char* __restrict__ mask_out) {
  __shared__ float s_mean[32], s_rstd[32];
  const int n = blockIdx.y;
  const int cpg = Ctot >> 5;
  if (threadIdx.x < 32) {
    float m, r;
    gn_group_moments(st1, st2, s.C1, s.C2, rps, n, threadIdx.x, cpg, HW, eps, &m, &r);
    s_mean[threadIdx.x] = m;
    s_rstd[threadIdx.x] = r;
    if (blockIdx.x == 0) { mean_out[n * 32 + threadIdx.x] = m; rstd_out[n * 32 + threadIdx.x] = r; }
  }
  __syncthreads();
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V, prow = threadIdx.x / V, rows = blockDim.x / V;
  const int c = v * 8;
  float g[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float gm = __ldg(gamma + c + i), rs_ = s_rstd[(c + i) / cpg];
    g[i] = gm * rs_;
    b[i] = fmaf(-s_mean[(c + i) / cpg], g[i], __ldg(beta + c + i));
  }
  const float keep_scale = drop_p > 0.f ? 1.0f / (1.0f - drop_p) : 1.0f;
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 4 * step) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix0 + u * step < HW) raw[u] = ldg_nc_u4(src_ptr(s, (long long)n * HW + pix0 + u * step, c));
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pix = pix0 + u * step;
      if (pix >= HW) continue;
      const long long gp = (long long)n * HW + pix;
      float x[8];
      unpack8(raw[u], x);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float z = fmaf(x[i], g[i], b[i]);
        x[i] = act ? silu_f(z) : z;
      }
      if (drop_p > 0.f) {
        bool keep[8];
        dropout_keep8(seed, (unsigned long long)gp * Ctot + c, drop_p, keep);
        unsigned bits = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { x[i] = keep[i] ? x[i] * keep_scale : 0.f; bits |= keep[i] ? (1u << i) : 0u; }
        if (mask_out) mask_out[gp * V + v] = (unsigned char)bits;
      }
      store8_fmt(y + gp * Ctot + c, x, FMT_F16);
    }
  }
}

// dx = dyh * rg + x * pp + qq (+ add_a + add_b); ab[n][c] = (sum dyh, sum dyh * x)
template <bool CS>
__global__ void __launch_bounds__(256, CS ? 2 : 3)
gn_bwd_apply2_kernel(GnSrc s, const __half* __restrict__ dyh, int HW, int Ctot, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ ab,
                     GnBwdOut o) {
  __shared__ float s1[32], s2[32];
  extern __shared__ float shc[];   // [C1] when o.colsum
  const int n = blockIdx.y;
  const int cpg = Ctot >> 5;
  if (threadIdx.x < 32) {
    const float rs_ = __ldg(rstd + n * 32 + threadIdx.x), mu_ = __ldg(mean + n * 32 + threadIdx.x);
    float a = 0.f, b = 0.f;
    for (int i = 0; i < cpg; ++i) {
      const int ch = threadIdx.x * cpg + i;
      const float gm = __ldg(gamma + ch);
      const float A = ab[((long long)n * Ctot + ch) * 2 + 0], B = ab[((long long)n * Ctot + ch) * 2 + 1];
      a = fmaf(gm, A, a);
      b = fmaf(gm, rs_ * (B - mu_ * A), b);          // sum dyh * xhat = rstd * (sum dyh * x - mean * sum dyh)
    }
    const float inv_m = 1.0f / ((float)cpg * (float)HW);
    s1[threadIdx.x] = a * inv_m;
    s2[threadIdx.x] = b * inv_m;
  }
  __syncthreads();
  const int V = Ctot >> 3;
  const int v = threadIdx.x % V, prow = threadIdx.x / V, rows = blockDim.x / V;
  const int c = v * 8;
  float rg[8], pp[8], qq[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int gi = (c + i) / cpg;
    const float rs_ = __ldg(rstd + n * 32 + gi), mu_ = __ldg(mean + n * 32 + gi);
    rg[i] = rs_ * __ldg(gamma + c + i);
    pp[i] = -rs_ * rs_ * s2[gi];
    qq[i] = -mu_ * pp[i] - rs_ * s1[gi];
  }
  const bool first = c < s.C1;
  const bool has_a = o.add_a && (first || o.add_a_full);
  const bool has_b = o.add_b && first;
  float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float amax = 0.f;
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 2 * step) {
    uint4 rx[2], rd[2], ra[2], rb[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (pix0 + u * step < HW) {
        const long long gp = (long long)n * HW + pix0 + u * step;
        rx[u] = ldg_nc_u4(src_ptr(s, gp, c));
        rd[u] = ldg_nc_u4(dyh + gp * Ctot + c);
        if (has_a) ra[u] = ldg_nc_u4(o.add_a + gp * o.lda + c);
        if (has_b) rb[u] = ldg_nc_u4(o.add_b + gp * o.ldb + c);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int pix = pix0 + u * step;
      if (pix >= HW) continue;
      const long long gp = (long long)n * HW + pix;
      float x[8], d[8];
      unpack8(rx[u], x);
      unpack8(rd[u], d);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        amax = fmaxf(amax, fabsf(d[i]));
        x[i] = fmaf(d[i], rg[i], fmaf(x[i], pp[i], qq[i]));
      }
      if (has_a) { float t[8]; unpack8(ra[u], t);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] += t[i]; }
      if (has_b) { float t[8]; unpack8(rb[u], t);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] += t[i]; }
#pragma unroll
      for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(x[i]));
      if (first) {
        store8(o.dx1 + gp * o.ldd1 + c, x);
        if (CS) {
#pragma unroll
          for (int i = 0; i < 8; ++i) cs[i] += x[i];
        }
      } else {
        store8(o.dx2 + gp * o.ldd2 + (c - s.C1), x);
      }
    }
  }
  if (o.overflow != nullptr && !(amax < 65504.f)) atomicAdd(o.overflow, 1);   // saturated (or NaN) fp16 gradient seen
  if (CS) gn_colsum_flush(o, shc, cs, s.C1, c, n, first);
}

// dgamma[c] += sum_n rstd * (ab1 - mean * ab0), dbeta[c] += sum_n ab0.  block = 32 channels x 8 sample lanes over the
// sample slice of blockIdx.y, four samples per lane in flight (one sample per iteration on one block column walked 32
// dependent load round trips: 20 us per call, 51 calls per step).
__global__ void __launch_bounds__(256)
gn_param_grad2_kernel(const float* __restrict__ ab, const float* __restrict__ mean, const float* __restrict__ rstd, int N,
                      int Ctot, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float sa[8][33], sb[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  const int cpg = Ctot >> 5;
  const int per = (N + gridDim.y - 1) / gridDim.y;
  const int n_lo = blockIdx.y * per, n_hi = min(N, n_lo + per);
  float a = 0.f, b = 0.f;
  if (c < Ctot) {
    const int g = c / cpg;
    for (int n0 = n_lo + ty; n0 < n_hi; n0 += 32) {
      float2 t[4];
      float rs_[4], mu_[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int n = n0 + 8 * u;
        const bool ok = n < n_hi;
        t[u] = ok ? __ldg(reinterpret_cast<const float2*>(ab + ((long long)n * Ctot + c) * 2)) : make_float2(0.f, 0.f);
        rs_[u] = ok ? __ldg(rstd + n * 32 + g) : 0.f;
        mu_[u] = ok ? __ldg(mean + n * 32 + g) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        a += t[u].x;
        b = fmaf(rs_[u], t[u].y - mu_[u] * t[u].x, b);
      }
    }
  }
  sa[ty][tx] = a; sb[ty][tx] = b;
  __syncthreads();
  if (ty == 0 && c < Ctot) {
    float ta = 0.f, tb = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { ta += sa[i][tx]; tb += sb[i][tx]; }
    atomicAdd(dbeta + c, ta);
    atomicAdd(dgamma + c, tb);
  }
}

// Cluster size and pixels per CTA for the single-pass kernels, or 0 for the two-kernel path.  Measured on the train64
// step (N = 256; profiles/r1_groupnorm_fused.md): the single pass wins only while a sample fits one or two CTAs of
// <= 72 KB (HW = 256: 0.021 vs 0.033 ms forward, 0.060 vs 0.078 ms backward at C = 128); with 4 - 8 CTAs per sample the
// load -> reduce -> barrier -> apply phases of a CTA serialise and only 1 - 3 CTAs fit an SM, so the kernel becomes
// latency bound (HW = 4096, C = 128: 2.03 vs 0.90 ms) although it moves 40 % fewer bytes.  B2D_GN_FUSED=0 disables the
// single pass, =2 forces it wherever the tile fits 200 KB at cluster 8 (the A/B switch behind those numbers).
static int gn_fused_plan(int HW, int Ctot, int* ppc, size_t* smem) {
  static const int mode = [] { const char* e = getenv("B2D_GN_FUSED"); return e ? atoi(e) : 1; }();
  if (mode == 0) return 0;
  const size_t extra = (size_t)(2 * Ctot + 64) * sizeof(float);
  const int max_cl = mode == 2 ? 8 : 2;
  for (int cl = 1; cl <= max_cl; cl *= 2) {
    const int p = (HW + cl - 1) / cl;
    const size_t bytes = (size_t)p * Ctot * 2 + extra;
    if (bytes <= (size_t)(cl == 8 && mode == 2 ? 200 : 72) * 1024) { *ppc = p; *smem = bytes; return cl; }
  }
  return 0;
}

template <typename... Args>
static cudaError_t gn_launch_cluster(void (*kern)(Args...), int cl, int N, int threads, size_t smem, cudaStream_t st,
                                     Args... args) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 4096);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)cl, (unsigned)N);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}

// blocks_per_sm: grid size in blocks per SM (more blocks = shorter pixel runs per block but better balance)
static int gn_launch_dims(int N, int HW, int Ctot, int blocks_per_sm, dim3* grid, int* threads) {
  const int V = Ctot / 8;
  int rows = 256 / V;
  if (rows < 1) rows = 1;
  *threads = V * rows;
  long long chunks = (HW + rows - 1) / rows;
  long long cap = ((long long)num_sms() * blocks_per_sm + N - 1) / N;
  if (chunks > cap) chunks = cap;
  if (chunks < 1) chunks = 1;
  *grid = dim3((unsigned)chunks, (unsigned)N);
  return 0;
}

static int gn_check(const b2d_gn_args* a, const char* who) {
  B2D_REQUIRE(a && a->x1, B2D_ERR_SHAPE, "%s: null input", who);
  const int C2 = a->x2 ? a->C2 : 0;
  const int Ctot = a->C1 + C2;
  B2D_REQUIRE(a->C1 % 8 == 0 && C2 % 8 == 0 && Ctot % 32 == 0 && Ctot <= 2048, B2D_ERR_SHAPE,
              "%s: channels (%d,%d) unsupported", who, a->C1, C2);
  B2D_REQUIRE(a->N > 0 && a->N <= 65535 && a->HW > 0, B2D_ERR_SHAPE, "%s: bad N=%d HW=%d", who, a->N, a->HW);
  B2D_REQUIRE(a->ld1 % 8 == 0 && (C2 == 0 || a->ld2 % 8 == 0) && aligned16(a->x1) && (C2 == 0 || aligned16(a->x2)),
              B2D_ERR_ALIGN, "%s: inputs must be 16-byte aligned rows", who);
  B2D_REQUIRE(a->drop_p >= 0.f && a->drop_p < 1.f, B2D_ERR_SHAPE, "%s: dropout p=%f", who, a->drop_p);
  return B2D_OK;
}

static GnSrc make_src(const b2d_gn_args* a) {
  GnSrc s;
  s.x1 = (const __half*)a->x1; s.ld1 = a->ld1; s.C1 = a->C1;
  s.x2 = (const __half*)a->x2; s.ld2 = a->ld2; s.C2 = a->x2 ? a->C2 : 0;
  return s;
}

}  // namespace b2d

using namespace b2d;

extern "C" size_t b2d_groupnorm_workspace_bytes(int N, int Ctot) {
  size_t a = (size_t)N * 32 * 2 * sizeof(double);
  size_t b = (size_t)N * Ctot * 2 * sizeof(float);
  return a > b ? a : b;
}

extern "C" int b2d_groupnorm_fwd(const b2d_gn_args* a, void* y, float* mean, float* rstd, void* workspace,
                                 size_t workspace_bytes, void* stream) {
  int rc = gn_check(a, "groupnorm_fwd");
  if (rc) return rc;
  B2D_REQUIRE(y && mean && rstd && workspace && a->gamma && a->beta, B2D_ERR_SHAPE, "groupnorm_fwd: null pointer");
  const GnSrc s = make_src(a);
  const int Ctot = s.C1 + s.C2;
  B2D_REQUIRE(workspace_bytes >= (size_t)a->N * 64 * sizeof(double), B2D_ERR_WORKSPACE, "groupnorm_fwd: workspace");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid; int threads;
  int ppc = 0; size_t fsmem = 0;
  if (const int cl = gn_fused_plan(a->HW, Ctot, &ppc, &fsmem)) {
    gn_launch_dims(a->N, a->HW, Ctot, 1, &grid, &threads);
    B2D_CHECK_CUDA(gn_launch_cluster(gn_fwd_fused_kernel, cl, a->N, threads, fsmem, st, s, a->HW, Ctot, ppc, a->gamma,
                                     a->beta, a->eps, a->act, a->drop_p, (unsigned long long)a->seed, (__half*)y, mean,
                                     rstd));
    B2D_CHECK_LAUNCH();
    return B2D_OK;
  }
  B2D_CHECK_CUDA(cudaMemsetAsync(workspace, 0, (size_t)a->N * 64 * sizeof(double), st));
  gn_launch_dims(a->N, a->HW, Ctot, 3, &grid, &threads);   // measured on the train64 step: 3 -> 3.84 ms, 8 -> 4.43 ms
  gn_stats_kernel<<<grid, threads, 2 * Ctot * sizeof(float), st>>>(s, a->HW, Ctot, (double*)workspace);
  B2D_CHECK_LAUNCH();
  gn_apply_kernel<<<grid, threads, 0, st>>>(s, a->HW, Ctot, (const double*)workspace, a->gamma, a->beta, a->eps,
                                            a->act, a->drop_p, a->seed, (__half*)y, mean, rstd);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

// ---- allocation-free variant (no counterpart in the reference) -------
// The two-kernel GroupNorm forward needs N * 64 doubles of zeroed scratch (per-(sample, group) sum / sum of squares).
// This variant keeps that scratch in a file-scope device array instead of a caller buffer, so a caller that may not
// allocate (under CUDA graph capture, in a serving loop) can run it.  Every GPU that loads the library holds the
// B2D_GN_STATIC_MAX_N * 512 bytes until exit.  All calls share the array: it is cleared on the stream before use, and
// two calls that overlap on the GPU (different streams, no ordering) race on it.
static __device__ __align__(256) unsigned char g_gn_static_scratch[(size_t)B2D_GN_STATIC_MAX_N * 64 * sizeof(double)];

// The array's name in host code is a host-side stand-in, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, taken once per device with that device current.
static int gn_static_scratch_ptr(void** out) {
  static std::once_flag once[16];
  static cudaError_t err[16];
  static void* ptr[16];
  int dev = 0;
  B2D_CHECK_CUDA(cudaGetDevice(&dev));
  B2D_REQUIRE(dev >= 0 && dev < 16, B2D_ERR_SHAPE, "groupnorm_fwd_static: device index %d not in [0, 16)", dev);
  std::call_once(once[dev], [&] { err[dev] = cudaGetSymbolAddress(&ptr[dev], g_gn_static_scratch); });
  B2D_CHECK_CUDA(err[dev]);
  *out = ptr[dev];
  return B2D_OK;
}

extern "C" int b2d_groupnorm_fwd_static(const b2d_gn_args* a, void* y, float* mean, float* rstd, void* stream) {
  B2D_REQUIRE(a, B2D_ERR_SHAPE, "groupnorm_fwd_static: null arguments");
  B2D_REQUIRE(a->N >= 1 && a->N <= B2D_GN_STATIC_MAX_N, B2D_ERR_WORKSPACE,
              "groupnorm_fwd_static: N = %d samples exceeds the built-in scratch (B2D_GN_STATIC_MAX_N = %d); use "
              "b2d_groupnorm_fwd with a caller workspace", a->N, B2D_GN_STATIC_MAX_N);
  void* ws = nullptr;
  int rc = gn_static_scratch_ptr(&ws);
  if (rc) return rc;
  return b2d_groupnorm_fwd(a, y, mean, rstd, ws, sizeof(g_gn_static_scratch), stream);
}

extern "C" int b2d_groupnorm_bwd(const b2d_gn_args* a, const void* dy, const float* mean, const float* rstd,
                                 void* dx1, int64_t lddx1, void* dx2, int64_t lddx2, const void* add_a, int64_t lda,
                                 const void* add_b, int64_t ldb, int add_a_full, float* dx1_colsum, int64_t ld_colsum,
                                 float* dgamma, float* dbeta, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = gn_check(a, "groupnorm_bwd");
  if (rc) return rc;
  const GnSrc s = make_src(a);
  const int Ctot = s.C1 + s.C2;
  B2D_REQUIRE(dy && mean && rstd && dx1 && workspace && a->gamma && a->beta && dgamma && dbeta, B2D_ERR_SHAPE,
              "groupnorm_bwd: null pointer");
  B2D_REQUIRE(s.C2 == 0 || dx2, B2D_ERR_SHAPE, "groupnorm_bwd: dx2 required for a two-source input");
  B2D_REQUIRE(workspace_bytes >= (size_t)a->N * Ctot * 2 * sizeof(float), B2D_ERR_WORKSPACE, "groupnorm_bwd: workspace");
  cudaStream_t st = (cudaStream_t)stream;
  B2D_CHECK_CUDA(cudaMemsetAsync(workspace, 0, (size_t)a->N * Ctot * 2 * sizeof(float), st));
  dim3 grid; int threads;
  GnBwdOut o;
  o.dx1 = (__half*)dx1; o.ldd1 = lddx1; o.dx2 = (__half*)dx2; o.ldd2 = lddx2;
  o.add_a = (const __half*)add_a; o.lda = lda; o.add_b = (const __half*)add_b; o.ldb = ldb; o.add_a_full = add_a_full;
  o.colsum = dx1_colsum; o.ldcs = ld_colsum; o.overflow = a->overflow;
  int ppc = 0; size_t fsmem = 0;
  if (const int cl = gn_fused_plan(a->HW, Ctot, &ppc, &fsmem)) {
    gn_launch_dims(a->N, a->HW, Ctot, 1, &grid, &threads);
    B2D_CHECK_CUDA(gn_launch_cluster(dx1_colsum ? gn_bwd_fused_kernel<true> : gn_bwd_fused_kernel<false>, cl, a->N, threads,
                                     fsmem, st, s, (const __half*)dy, a->HW, Ctot,
                                     ppc, mean, rstd, a->gamma, a->beta, a->act, a->drop_p,
                                     (unsigned long long)a->seed, (float*)workspace, o));
    B2D_CHECK_LAUNCH();
    gn_param_grad_kernel<<<(Ctot + 31) / 32, 256, 0, st>>>((const float*)workspace, a->N, Ctot, dgamma, dbeta);
    B2D_CHECK_LAUNCH();
    return B2D_OK;
  }
  gn_launch_dims(a->N, a->HW, Ctot, 8, &grid, &threads);   // measured: 8 -> 10.4 ms, 3 -> 12.1 ms
  gn_bwd_reduce_kernel<<<grid, threads, 2 * Ctot * sizeof(float), st>>>(
      s, (const __half*)dy, a->HW, Ctot, mean, rstd, a->gamma, a->beta, a->act, a->drop_p, a->seed,
      (float*)workspace);
  B2D_CHECK_LAUNCH();
  if (dx1_colsum)
    gn_bwd_apply_kernel<true><<<grid, threads, s.C1 * sizeof(float), st>>>(
        s, (const __half*)dy, a->HW, Ctot, mean, rstd, a->gamma, a->beta, a->act, a->drop_p, a->seed,
        (const float*)workspace, o);
  else
    gn_bwd_apply_kernel<false><<<grid, threads, 0, st>>>(
        s, (const __half*)dy, a->HW, Ctot, mean, rstd, a->gamma, a->beta, a->act, a->drop_p, a->seed,
        (const float*)workspace, o);
  B2D_CHECK_LAUNCH();
  gn_param_grad_kernel<<<(Ctot + 31) / 32, 256, 0, st>>>((const float*)workspace, a->N, Ctot, dgamma, dbeta);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_pair_stats(const void* x, int64_t ld, int C, int N, int HW, double* st, void* stream) {
  B2D_REQUIRE(x && st && C > 0 && C % 8 == 0 && C <= 2048 && N > 0 && N <= 65535 && HW > 0, B2D_ERR_SHAPE,
              "pair_stats: bad arguments (C=%d N=%d HW=%d)", C, N, HW);
  B2D_REQUIRE(ld % 8 == 0 && aligned16(x), B2D_ERR_ALIGN, "pair_stats: input must have 16-byte aligned rows");
  dim3 grid; int threads;
  gn_launch_dims(N, HW, C, 3, &grid, &threads);
  pair_stats_kernel<<<grid, threads, C * sizeof(float), (cudaStream_t)stream>>>((const __half*)x, ld, C, HW, st);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_groupnorm_apply(const b2d_gn_args* a, const double* st1, const double* st2, int rows_per_sample,
                                   void* y, float* mean, float* rstd, uint8_t* mask_out, void* stream) {
  int rc = gn_check(a, "groupnorm_apply");
  if (rc) return rc;
  const GnSrc s = make_src(a);
  const int Ctot = s.C1 + s.C2;
  B2D_REQUIRE(y && mean && rstd && st1 && (s.C2 == 0 || st2) && a->gamma && a->beta && rows_per_sample >= 1, B2D_ERR_SHAPE,
              "groupnorm_apply: null pointer");
  B2D_REQUIRE(s.C1 % 2 == 0 && s.C2 % 2 == 0 && (Ctot / 32) % 2 == 0, B2D_ERR_SHAPE,
              "groupnorm_apply: groups of %d channels do not split into channel pairs", Ctot / 32);
  dim3 grid; int threads;
  // CTAs per SM's worth of grid: 2-3 measured best (scripts/bench_gn_fused.py: forward class 2.16 -> 2.0 ms, backward
  // 5.17 -> 4.7 ms against 8): fewer, longer CTAs amortise the per-CTA moment / constant set-up
  static const int bps_f = [] { const char* e = getenv("B2D_GN_BPS_F"); return e ? atoi(e) : 3; }();
  gn_launch_dims(a->N, a->HW, Ctot, bps_f, &grid, &threads);
  gn_apply2_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(s, a->HW, Ctot, st1, st2, rows_per_sample, a->gamma, a->beta,
                                                               a->eps, a->act, a->drop_p, a->seed, (__half*)y, mean, rstd,
                                                               mask_out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_groupnorm_bwd_apply(const b2d_gn_args* a, const void* dyh, const float* mean, const float* rstd,
                                       const float* ab, void* dx1, int64_t lddx1, void* dx2, int64_t lddx2,
                                       const void* add_a, int64_t lda, const void* add_b, int64_t ldb, int add_a_full,
                                       float* dx1_colsum, int64_t ld_colsum, float* dgamma, float* dbeta, void* stream) {
  int rc = gn_check(a, "groupnorm_bwd_apply");
  if (rc) return rc;
  const GnSrc s = make_src(a);
  const int Ctot = s.C1 + s.C2;
  B2D_REQUIRE(dyh && mean && rstd && ab && dx1 && a->gamma && dgamma && dbeta, B2D_ERR_SHAPE,
              "groupnorm_bwd_apply: null pointer");
  B2D_REQUIRE(s.C2 == 0 || dx2, B2D_ERR_SHAPE, "groupnorm_bwd_apply: dx2 required for a two-source input");
  cudaStream_t st = (cudaStream_t)stream;
  GnBwdOut o;
  o.dx1 = (__half*)dx1; o.ldd1 = lddx1; o.dx2 = (__half*)dx2; o.ldd2 = lddx2;
  o.add_a = (const __half*)add_a; o.lda = lda; o.add_b = (const __half*)add_b; o.ldb = ldb; o.add_a_full = add_a_full;
  o.colsum = dx1_colsum; o.ldcs = ld_colsum; o.overflow = a->overflow;
  dim3 grid; int threads;
  static const int bps_b = [] { const char* e = getenv("B2D_GN_BPS_B"); return e ? atoi(e) : 3; }();
  gn_launch_dims(a->N, a->HW, Ctot, bps_b, &grid, &threads);
  if (dx1_colsum)
    gn_bwd_apply2_kernel<true><<<grid, threads, s.C1 * sizeof(float), st>>>(s, (const __half*)dyh, a->HW, Ctot, mean, rstd,
                                                                           a->gamma, ab, o);
  else
    gn_bwd_apply2_kernel<false><<<grid, threads, 0, st>>>(s, (const __half*)dyh, a->HW, Ctot, mean, rstd, a->gamma, ab, o);
  B2D_CHECK_LAUNCH();
  gn_param_grad2_kernel<<<dim3((unsigned)((Ctot + 31) / 32), (unsigned)((a->N + 31) / 32 < 8 ? (a->N + 31) / 32 : 8)), 256, 0, st>>>(
      ab, mean, rstd, a->N, Ctot, dgamma, dbeta);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
