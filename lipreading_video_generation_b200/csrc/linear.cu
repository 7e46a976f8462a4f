// linear.cu -- small fp32 Linear layers of the conditioning path, forward and backward:
// time_embed (unet.py:483-487), the 17 ResBlock emb_layers SiLU->Linear (unet.py:211-217) batched
// into one GEMM, and AudioFeatureTransformer Linear+ReLU (unet_audio.py:21-30).
// M = images in the batch (<= a few hundred), K, N <= a few thousand: ~0.002 % of the step's FLOPs,
// so a plain fp32 SIMT tiled GEMM (exact fp32 parity with the reference) is the right tool here;
// the tensor-core kernels are reserved for the convolutions and attention.
#include "common.cuh"

namespace b2d {

// C[m][n] (=|+=) epi( sum_k fa(A(m,k)) * fb(B(k,n)) )
//   A(m,k) = a[m*sam + k*sak] (* (amask[..] > 0) when amask),  B(k,n) = b[k*sbk + n*sbn] or 1 when b == null
//   fa / fb: optional SiLU.  epi: + bias[n], ReLU, * silu'(aux[m*ldc + n]).
struct GemmF32 {
  const float* a; long long sam, sak; const float* amask; int act_a;
  const float* b; long long sbk, sbn; int act_b;
  float* c; long long ldc;
  const float* bias; int relu_out; const float* dsilu_aux; int accumulate;
  int M, N, K;
  int ksplit;   // > 1: gridDim.z CTAs share K, partial sums are added to c with atomics (c holds zeros or the value to
                // accumulate onto); only for epilogues that are linear in the sum (no bias, no ReLU)
};

__device__ __forceinline__ float silu_grad(float z) {
  const float sg = 1.0f / (1.0f + expf(-z));
  return sg * (1.0f + z * (1.0f - sg));
}
__device__ __forceinline__ float silu_exact(float z) { return z / (1.0f + expf(-z)); }

__global__ void __launch_bounds__(256)
gemm_f32_kernel(GemmF32 g) {
  __shared__ float As[32][33];
  __shared__ float Bs[32][33];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // 16 x 16 threads, 2x2 outputs each
  const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
  int k_lo = 0, k_hi = g.K;
  if (g.ksplit > 1) {
    const int per = ((g.K + g.ksplit - 1) / g.ksplit + 31) & ~31;
    k_lo = blockIdx.z * per;
    k_hi = min(g.K, k_lo + per);
  }
  for (int k0 = k_lo; k0 < k_hi; k0 += 32) {
    // consecutive threads walk whichever index of the operand is contiguous in memory (the transposed operands of
    // the forward (w[n][k]) and of the weight gradient (dy[m][n] as A(n, m)) were read with a stride of a full row)
    const bool a_k_fast = g.sak == 1, b_n_fast = (g.sbn == 1) || !g.b;
    for (int i = threadIdx.x; i < 1024; i += 256) {
      const int r = i >> 5, cidx = i & 31;
      {
        const int mi = a_k_fast ? r : cidx, ki = a_k_fast ? cidx : r;
        const int m = m0 + mi, k = k0 + ki;
        float v = 0.f;
        if (m < g.M && k < k_hi) {
          const long long off = (long long)m * g.sam + (long long)k * g.sak;
          v = g.a[off];
          if (g.amask) v = g.amask[off] > 0.f ? v : 0.f;
          if (g.act_a) v = silu_exact(v);
        }
        As[mi][ki] = v;
      }
      {
        const int ki = b_n_fast ? r : cidx, ni = b_n_fast ? cidx : r;
        const int k = k0 + ki, n = n0 + ni;
        float v = 0.f;
        if (k < k_hi && n < g.N) {
          v = g.b ? g.b[(long long)k * g.sbk + (long long)n * g.sbn] : 1.0f;
          if (g.act_b) v = silu_exact(v);
        }
        Bs[ki][ni] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const float a0 = As[ty][k], a1 = As[ty + 16][k];
      const float b0 = Bs[k][tx], b1 = Bs[k][tx + 16];
      acc[0][0] = fmaf(a0, b0, acc[0][0]); acc[0][1] = fmaf(a0, b1, acc[0][1]);
      acc[1][0] = fmaf(a1, b0, acc[1][0]); acc[1][1] = fmaf(a1, b1, acc[1][1]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int m = m0 + ty + 16 * i, n = n0 + tx + 16 * j;
      if (m < g.M && n < g.N) {
        float v = acc[i][j];
        if (g.bias) v += g.bias[n];
        if (g.relu_out) v = v > 0.f ? v : 0.f;
        const long long off = (long long)m * g.ldc + n;
        if (g.dsilu_aux) v *= silu_grad(g.dsilu_aux[off]);
        if (g.ksplit > 1) atomicAdd(&g.c[off], v);
        else g.c[off] = g.accumulate ? g.c[off] + v : v;
      }
    }
}

static int launch_gemm(const GemmF32& g_in, void* stream) {
  GemmF32 g = g_in;
  if (g.M <= 0 || g.N <= 0) return B2D_OK;
  dim3 grid((unsigned)((g.N + 31) / 32), (unsigned)((g.M + 31) / 32));
  // a long K over a few output tiles (dx of the batched emb_layers: 256 x 256 outputs, K = 2816 -> 64 CTAs walking 88
  // tiles each, 0.27 ms) is split over gridDim.z; needs an epilogue that is linear in the sum and a contiguous c
  g.ksplit = 1;
  const long long ctas = (long long)grid.x * grid.y;
  if (!g.bias && !g.relu_out && g.K >= 512 && ctas * 2 <= num_sms() && g.ldc == g.N) {
    int ks = (int)((2LL * num_sms() + ctas - 1) / ctas);
    if (ks > g.K / 128) ks = g.K / 128;
    if (ks > 1) {
      g.ksplit = ks;
      grid.z = (unsigned)ks;
      if (!g.accumulate)
        B2D_CHECK_CUDA(cudaMemsetAsync(g.c, 0, (size_t)g.M * g.N * sizeof(float), (cudaStream_t)stream));
    }
  }
  gemm_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

}  // namespace b2d

using namespace b2d;

// y[M][Nout] = act_out( f_in(x)[M][K] @ w[Nout][K]^T + bias ),  f_in = SiLU when silu_in, act_out = ReLU when relu_out
extern "C" int b2d_linear_fwd(const float* x, const float* w, const float* bias, float* y, int M, int K, int Nout,
                              int silu_in, int relu_out, void* stream) {
  B2D_REQUIRE(x && w && y && M > 0 && K > 0 && Nout > 0, B2D_ERR_SHAPE, "linear_fwd: bad arguments");
  GemmF32 g{};
  g.a = x; g.sam = K; g.sak = 1; g.act_a = silu_in;
  g.b = w; g.sbk = 1; g.sbn = K;
  g.c = y; g.ldc = Nout; g.bias = bias; g.relu_out = relu_out;
  g.M = M; g.N = Nout; g.K = K;
  return launch_gemm(g, stream);
}

// Given dy (gradient w.r.t. y) and, when relu_out, y itself (mask y > 0):
//   dx[M][K]     = (dy*mask) @ w   (* silu'(x) when silu_in)        [written, or skipped when dx == NULL]
//   dw[Nout][K] += (dy*mask)^T @ f_in(x)
//   db[Nout]    += column sums of dy*mask                            [skipped when db == NULL]
extern "C" int b2d_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dx, float* dw,
                              float* db, int M, int K, int Nout, int silu_in, int relu_out, void* stream) {
  B2D_REQUIRE(x && w && dy && dw && (!relu_out || y), B2D_ERR_SHAPE, "linear_bwd: bad arguments");
  const float* mask = relu_out ? y : nullptr;
  int rc;
  if (dx) {
    GemmF32 g{};
    g.a = dy; g.sam = Nout; g.sak = 1; g.amask = mask;
    g.b = w; g.sbk = K; g.sbn = 1;
    g.c = dx; g.ldc = K; g.dsilu_aux = silu_in ? x : nullptr;
    g.M = M; g.N = K; g.K = Nout;
    if ((rc = launch_gemm(g, stream))) return rc;
  }
  {
    GemmF32 g{};
    g.a = dy; g.sam = 1; g.sak = Nout; g.amask = mask;   // A(m'=n, k'=m) = dy[m][n]
    g.b = x; g.sbk = K; g.sbn = 1; g.act_b = silu_in;    // B(k'=m, n'=k) = f_in(x[m][k])
    g.c = dw; g.ldc = K; g.accumulate = 1;
    g.M = Nout; g.N = K; g.K = M;
    if ((rc = launch_gemm(g, stream))) return rc;
  }
  if (db) {
    GemmF32 g{};
    g.a = dy; g.sam = 1; g.sak = Nout; g.amask = mask;
    g.b = nullptr;
    g.c = db; g.ldc = 1; g.accumulate = 1;
    g.M = Nout; g.N = 1; g.K = M;
    if ((rc = launch_gemm(g, stream))) return rc;
  }
  return B2D_OK;
}
