// elementwise.cu -- layout / gather / reduction helpers around the tensor-core kernels:
// conditioning assembly (unet_audio.py:53-61), nearest upsample (unet.py:112-122), stride-2 dgrad
// zero insertion, NCHW fp32 <-> pixel-major bf16, per-sample column sums (bias / emb gradients),
// sinusoidal timestep embedding (utils.py:140-158).  All HBM-bound, 16-byte vector accesses.
#include "common.cuh"

namespace b2d {

__device__ __forceinline__ void ld8(const __half* p, float* v) {
  uint4 u = ldg_nc_u4(p);
  float2 a = unpack_f16x2(u.x), b = unpack_f16x2(u.y), c = unpack_f16x2(u.z), d = unpack_f16x2(u.w);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
}
__device__ __forceinline__ void st8(__half* p, const float* v) {
  uint4 u;
  u.x = pack_f16x2(v[0], v[1]); u.y = pack_f16x2(v[2], v[3]);
  u.z = pack_f16x2(v[4], v[5]); u.w = pack_f16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// PyTorch 'nearest' source index: floor(dst * in / out) computed in float (F.interpolate default)
__device__ __forceinline__ int nearest_src(int dst, int in_size, int out_size) {
  if (in_size == out_size) return dst;
  const float scale = (float)in_size / (float)out_size;
  int s = (int)floorf((float)dst * scale);
  return s < in_size - 1 ? s : in_size - 1;
}

// ------------------------------------------------------------------ conditioning assembly
struct AssembleParams {
  const float* xt;     // (N, Cin, S, S)
  const float* cond;   // (N, Cc, Hc, Wc) or null
  const float* wcond;  // (Cim, Cc)
  const float* audio;  // (N, A)
  int N, S, Cin, Cc, Hc, Wc, Cim, A, Cpad;
};

// grid = (pixel chunks, N); a thread owns one 8-channel vector (fixed for its lifetime) and walks pixels: everything
// that depends only on the channel (class of each channel, the 1x1 cond_conv_in weights, the per-sample audio
// values) is hoisted out of the pixel loop.  Cc <= 4 (the reference's conditioning image has 3 channels).
__global__ void __launch_bounds__(256)
assemble_input_kernel(AssembleParams p, __half* __restrict__ out) {
  // Two roles (blockIdx.z): the vectors that hold x_t / cond_conv_in channels (9 of 32 at the train.py configuration)
  // need loads and ~50 FMAs per pixel, the others (audio values, zero padding) are a pure store stream.  With one
  // pixel per warp both paths ran one after the other in every warp, the store lanes idling behind the load
  // latency of the others (0.51 ms for 537 MB); split by role every warp is uniform.
  const int V = p.Cpad >> 3;
  const int vg = min(V, (p.Cin + p.Cim + 7) >> 3);
  const int nv = blockIdx.z == 0 ? vg : V - vg;
  if (nv == 0) return;
  const int rows = blockDim.x / nv;
  if ((int)threadIdx.x >= rows * nv) return;
  const int v = (blockIdx.z == 0 ? 0 : vg) + threadIdx.x % nv;
  const int prow = threadIdx.x / nv;
  const int n = blockIdx.y;
  const int HW = p.S * p.S;
  int kind[8];            // 0 zero, 1 x_t channel, 2 cond_conv_in output, 3 audio
  float wq[8][4];         // kind 2: weights; kind 3: wq[j][0] = the value itself
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = v * 8 + j;
    kind[j] = 0;
    wq[j][0] = wq[j][1] = wq[j][2] = wq[j][3] = 0.f;
    if (c < p.Cin) kind[j] = 1;
    else if (c < p.Cin + p.Cim) {
      kind[j] = 2;
#pragma unroll
      for (int k = 0; k < 4; ++k)                     // fixed trip count: a runtime bound put wq in local memory
        if (k < p.Cc) wq[j][k] = __ldg(p.wcond + (long long)(c - p.Cin) * p.Cc + k);
    } else if (c < p.Cin + p.Cim + p.A) {
      kind[j] = 3;
      wq[j][0] = __ldg(p.audio + (long long)n * p.A + (c - p.Cin - p.Cim));
    }
  }
  bool any_cond = false, any_xt = false;
#pragma unroll
  for (int j = 0; j < 8; ++j) { any_cond |= kind[j] == 2; any_xt |= kind[j] == 1; }
  const int step = gridDim.x * rows;
  __half* const obase = out + (long long)n * HW * p.Cpad + v * 8;
  if (!any_cond && !any_xt) {
    // audio / zero-padding vectors (2/3 of the tensor): the value does not depend on the pixel -- a pure store stream
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = kind[j] == 3 ? wq[j][0] : 0.f;
    uint4 u;
    u.x = pack_f16x2(o[0], o[1]); u.y = pack_f16x2(o[2], o[3]); u.z = pack_f16x2(o[4], o[5]); u.w = pack_f16x2(o[6], o[7]);
    for (int pix = blockIdx.x * rows + prow; pix < HW; pix += step)
      *reinterpret_cast<uint4*>(obase + (long long)pix * p.Cpad) = u;
    return;
  }
  // general vectors (x_t and cond_conv_in channels, 9 of 32 lanes): every channel is the same arithmetic form
  //   o_j = w_j0 cv0 + w_j1 cv1 + w_j2 cv2 + w_j3 cv3 + e_j x_j + f_j
  // with per-lane coefficients (x_t: e = 1; audio: f = value; zero padding: all 0), so the loop body has no per-channel
  // branches (the first version selected on kind[j] per element: ~310 instructions per pixel, issue bound at 0.98 ms).
  float ej[8], fj[8];
  long long xoff[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    ej[j] = kind[j] == 1 ? 1.f : 0.f;
    fj[j] = kind[j] == 3 ? wq[j][0] : 0.f;
    if (kind[j] != 2) { wq[j][0] = wq[j][1] = wq[j][2] = wq[j][3] = 0.f; }
    const int cx = (v * 8 + j < p.Cin) ? v * 8 + j : p.Cin - 1;       // clamped: loaded but multiplied by e = 0
    xoff[j] = ((long long)n * p.Cin + cx) * HW;
  }
  const long long cstride = (long long)p.Hc * p.Wc;
  const float* cbase = p.cond ? p.cond + (long long)n * p.Cc * cstride : p.xt;
  const int ncc = p.cond ? p.Cc : 0;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 2 * step) {
    float cv[2][4], xv[2][8];
#pragma unroll
    for (int u = 0; u < 2; ++u) {                     // the loads of two pixels first
      const int pix = pix0 + u * step < HW ? pix0 + u * step : pix0;
      const int h = pix / p.S, w = pix - h * p.S;
      const int so = nearest_src(h, p.Hc, p.S) * p.Wc + nearest_src(w, p.Wc, p.S);
#pragma unroll
      for (int k = 0; k < 4; ++k) cv[u][k] = k < ncc ? __ldg(cbase + k * cstride + so) : 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) xv[u][j] = any_xt ? __ldg(p.xt + xoff[j] + pix) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int pix = pix0 + u * step;
      if (pix >= HW) continue;
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        o[j] = fmaf(wq[j][3], cv[u][3], fmaf(wq[j][2], cv[u][2], fmaf(wq[j][1], cv[u][1],
               fmaf(wq[j][0], cv[u][0], fmaf(ej[j], xv[u][j], fj[j])))));
      store8_fmt(obase + (long long)pix * p.Cpad, o, FMT_F16);   // forward tensor
    }
  }
}

// backward: d_wcond[c][k] += sum_{n,pix} dX[n,pix,Cin+c] * cond[n,k,src(pix)];  d_audio[n][a] += sum_pix dX[n,pix,Cin+Cim+a]
__global__ void __launch_bounds__(256)
assemble_input_bwd_kernel(AssembleParams p, const __half* __restrict__ dX, float* __restrict__ d_wcond,
                          float* __restrict__ d_audio) {
  extern __shared__ float sh[];  // [Cim*Cc + A]
  const int nsh = p.Cim * p.Cc + p.A;
  for (int i = threadIdx.x; i < nsh; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int n = blockIdx.y;
  const int V = (p.Cin + p.Cim + p.A + 7) >> 3;  // vectors that carry gradient
  const int v = threadIdx.x % V;
  const int rows = blockDim.x / V;
  const int prow = threadIdx.x / V;
  float acc[8][3];
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[j][0] = acc[j][1] = acc[j][2] = 0.f; }
  if (prow < rows) {
    // per-thread channel classes as multipliers (no per-element branches), four pixels in flight per thread (one
    // dependent 16-byte load per iteration left the kernel at 1.4 TB/s: 0.30 ms for 403 MB)
    float mc[8], ma[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = v * 8 + j;
      mc[j] = (c >= p.Cin && c < p.Cin + p.Cim) ? 1.f : 0.f;
      ma[j] = (c >= p.Cin + p.Cim) ? 1.f : 0.f;
    }
    const int HW = p.S * p.S, step = gridDim.x * rows;
    const long long cplane = (long long)p.Hc * p.Wc;
    const float* cbase = p.Cim > 0 ? p.cond + (long long)n * p.Cc * cplane : nullptr;
    for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 4 * step) {
      uint4 raw[4];
      float cv[4][3];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int pix = pix0 + u * step;
        cv[u][0] = cv[u][1] = cv[u][2] = 0.f;
        if (pix < HW) {
          raw[u] = ldg_nc_u4(dX + ((long long)n * HW + pix) * p.Cpad + v * 8);
          if (cbase) {
            const int h = pix / p.S, w = pix - h * p.S;
            const int so = nearest_src(h, p.Hc, p.S) * p.Wc + nearest_src(w, p.Wc, p.S);
#pragma unroll
            for (int k = 0; k < 3; ++k) if (k < p.Cc) cv[u][k] = __ldg(cbase + k * cplane + so);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (pix0 + u * step >= HW) continue;
        float d[8];
        unpack8(raw[u], d);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float dc = d[j] * mc[j];
          acc[j][0] = fmaf(dc, cv[u][0], fmaf(d[j], ma[j], acc[j][0]));
          acc[j][1] = fmaf(dc, cv[u][1], acc[j][1]);
          acc[j][2] = fmaf(dc, cv[u][2], acc[j][2]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = v * 8 + j;
      if (c >= p.Cin && c < p.Cin + p.Cim) {
        for (int k = 0; k < p.Cc && k < 3; ++k) atomicAdd(&sh[(c - p.Cin) * p.Cc + k], acc[j][k]);
      } else if (c >= p.Cin + p.Cim && c < p.Cin + p.Cim + p.A) {
        atomicAdd(&sh[p.Cim * p.Cc + (c - p.Cin - p.Cim)], acc[j][0]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < p.Cim * p.Cc; i += blockDim.x) atomicAdd(&d_wcond[i], sh[i]);
  for (int i = threadIdx.x; i < p.A; i += blockDim.x) atomicAdd(&d_audio[(long long)n * p.A + i], sh[p.Cim * p.Cc + i]);
}

// ------------------------------------------------------------------ nearest 2x upsample and its adjoint
__global__ void __launch_bounds__(256)
upsample2x_kernel(const __half* __restrict__ x, __half* __restrict__ y, int N, int H, int W, int C) {
  const int V = C >> 3;
  const long long total = (long long)N * 4 * H * W * V;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % V);
    long long r = i / V;
    const int wo = (int)(r % (2 * W)); r /= (2 * W);
    const int ho = (int)(r % (2 * H));
    const int n = (int)(r / (2 * H));
    const uint4 u = ldg_nc_u4(x + (((long long)n * H + (ho >> 1)) * W + (wo >> 1)) * C + v * 8);
    *reinterpret_cast<uint4*>(y + (((long long)n * 2 * H + ho) * 2 * W + wo) * C + v * 8) = u;
  }
}
__global__ void __launch_bounds__(256)
upsample2x_bwd_kernel(const __half* __restrict__ dy, __half* __restrict__ dx, int N, int H, int W, int C) {
  const int V = C >> 3;
  const long long total = (long long)N * H * W * V;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % V);
    long long r = i / V;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        float t[8];
        ld8(dy + (((long long)n * 2 * H + 2 * h + a) * 2 * W + 2 * w + b) * C + v * 8, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) s[j] += t[j];
      }
    st8(dx + (((long long)n * H + h) * W + w) * C + v * 8, s);
  }
}
// z[n, 2h, 2w] = dy[n, h, w], zero elsewhere: turns the stride-2 conv dgrad into a stride-1 conv
__global__ void __launch_bounds__(256)
zero_insert2x_kernel(const __half* __restrict__ dy, __half* __restrict__ z, int N, int Ho, int Wo, int C) {
  const int V = C >> 3;
  const long long total = (long long)N * 4 * Ho * Wo * V;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % V);
    long long r = i / V;
    const int w = (int)(r % (2 * Wo)); r /= (2 * Wo);
    const int h = (int)(r % (2 * Ho));
    const int n = (int)(r / (2 * Ho));
    uint4 u = make_uint4(0, 0, 0, 0);
    if (((h | w) & 1) == 0) u = ldg_nc_u4(dy + (((long long)n * Ho + (h >> 1)) * Wo + (w >> 1)) * C + v * 8);
    *reinterpret_cast<uint4*>(z + i * 8) = u;
  }
}

// fp32 (N, C, HW) -> bf16 (N, HW, Cpad), zero padded channels
__global__ void __launch_bounds__(256)
nchw_to_pixel_major_kernel(const float* __restrict__ src, __half* __restrict__ dst, int N, int C, int HW, int Cpad,
                           const float* __restrict__ scale_ptr) {
  const float sc = scale_ptr ? __ldg(scale_ptr) : 1.0f;
  const int V = Cpad >> 3;
  const long long total = (long long)N * HW * V;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % V);
    const long long gp = i / V;
    const int n = (int)(gp / HW);
    const int pix = (int)(gp - (long long)n * HW);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = v * 8 + j;
      o[j] = c < C ? src[((long long)n * C + c) * HW + pix] * sc : 0.f;
    }
    st8(dst + gp * Cpad + v * 8, o);
  }
}

// ------------------------------------------------------------------ weight packing (fp32 state_dict -> fp16 operand)
__global__ void __launch_bounds__(256)
pack_conv_weight_kernel(const float* __restrict__ w, long long s_o, long long s_i, long long s_k, int Cout, int Cin,
                        int K, int transpose, __half* __restrict__ out, int rows_pad, int inner_pad) {
  const long long total = (long long)rows_pad * K * inner_pad;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % inner_pad);
    const long long rk = i / inner_pad;
    const int k = (int)(rk % K);
    const int r = (int)(rk / K);
    const int co = transpose ? c : r, ci = transpose ? r : c;
    const int kw = transpose ? K - 1 - k : k;
    float v = 0.f;
    if (co < Cout && ci < Cin) v = fminf(fmaxf(__ldg(w + co * s_o + ci * s_i + kw * s_k), -65504.f), 65504.f);
    out[i] = __float2half_rn(v);
  }
}

// Every convolution's operand copy in ONE launch (the refresh after an optimiser step): a CTA owns PACK_BLOCK_ELEMS
// consecutive output elements of one descriptor, found by binary search over the descriptors' first-CTA index.
constexpr int PACK_BLOCK_ELEMS = 4096;
__global__ void __launch_bounds__(256)
pack_conv_weights_multi_kernel(const B2dPackDesc* __restrict__ descs, int n_descs) {
  __shared__ int s_idx;
  if (threadIdx.x == 0) {
    int lo = 0, hi = n_descs - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (descs[mid].block_begin <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
    }
    s_idx = lo;
  }
  __syncthreads();
  const B2dPackDesc d = descs[s_idx];
  const long long total = (long long)d.rows_pad * d.K * d.inner_pad;
  const long long base = (long long)((int)blockIdx.x - d.block_begin) * PACK_BLOCK_ELEMS;
  const long long end = base + PACK_BLOCK_ELEMS < total ? base + PACK_BLOCK_ELEMS : total;
  __half* __restrict__ out = reinterpret_cast<__half*>(d.out);
  if ((d.inner_pad & 7) == 0) {
    // 8 consecutive inner-index elements per thread: one index decomposition and one 16-byte store per 8 outputs
    // (the element-wise form spent its time in 64-bit div / mod: ncu IPC 2.6 at 0.95 TB/s)
    for (long long i = base + 8 * threadIdx.x; i < end; i += 8 * 256) {
      const int c = (int)(i % d.inner_pad);
      const long long rk = i / d.inner_pad;
      const int k = (int)(rk % d.K);
      const int r = (int)(rk / d.K);
      const int kw = d.transpose ? d.K - 1 - k : k;
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int co = d.transpose ? c + j : r, ci = d.transpose ? r : c + j;
        v[j] = (co < d.Cout && ci < d.Cin)
                   ? fminf(fmaxf(__ldg(d.w + co * d.s_o + ci * d.s_i + kw * d.s_k), -65504.f), 65504.f) : 0.f;
      }
      uint4 u;
      u.x = pack_f16x2_nosat(v[0], v[1]); u.y = pack_f16x2_nosat(v[2], v[3]);
      u.z = pack_f16x2_nosat(v[4], v[5]); u.w = pack_f16x2_nosat(v[6], v[7]);
      *reinterpret_cast<uint4*>(out + i) = u;
    }
    return;
  }
  for (long long i = base + threadIdx.x; i < end; i += 256) {
    const int c = (int)(i % d.inner_pad);
    const long long rk = i / d.inner_pad;
    const int k = (int)(rk % d.K);
    const int r = (int)(rk / d.K);
    const int co = d.transpose ? c : r, ci = d.transpose ? r : c;
    const int kw = d.transpose ? d.K - 1 - k : k;
    float v = 0.f;
    if (co < d.Cout && ci < d.Cin) v = fminf(fmaxf(__ldg(d.w + co * d.s_o + ci * d.s_i + kw * d.s_k), -65504.f), 65504.f);
    out[i] = __float2half_rn(v);
  }
}

// ------------------------------------------------------------------ per-sample column sums
// out[n][c] += sum_pix x[n, pix, c_off + c]   (emb_out gradient unet.py:266; bias grads = sum over n)
__global__ void __launch_bounds__(256)
colsum_kernel(const __half* __restrict__ x, long long ld, int c_off, int C, int HW, float* __restrict__ out,
              long long ldout) {
  extern __shared__ float sh[];  // [C]
  for (int i = threadIdx.x; i < C; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int n = blockIdx.y;
  const int V = C >> 3;
  const int v = threadIdx.x % V;
  const int rows = blockDim.x / V;
  const int prow = threadIdx.x / V;
  float a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int step = gridDim.x * rows;
  for (int pix0 = blockIdx.x * rows + prow; pix0 < HW; pix0 += 4 * step) {   // four 16-byte loads in flight
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (pix0 + u * step < HW) raw[u] = ldg_nc_u4(x + ((long long)n * HW + pix0 + u * step) * ld + c_off + v * 8);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (pix0 + u * step >= HW) continue;
      float t[8];
      unpack8(raw[u], t);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += t[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) atomicAdd(&sh[v * 8 + j], a[j]);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) atomicAdd(&out[(long long)n * ldout + i], sh[i]);
}

// out[c] += sum_r in[r][c]
// block = 32 columns x 8 row lanes; each lane sums every 8th row with four loads in flight
__global__ void __launch_bounds__(256)
reduce_rows_kernel(const float* __restrict__ in, int R, int C, long long ld, float* __restrict__ out) {
  __shared__ float sh[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float a = 0.f;
  if (c < C) {
    for (int r0 = ty; r0 < R; r0 += 32) {
      float t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) t[u] = (r0 + 8 * u < R) ? __ldg(in + (long long)(r0 + 8 * u) * ld + c) : 0.f;
      a += (t[0] + t[1]) + (t[2] + t[3]);
    }
  }
  sh[ty][tx] = a;
  __syncthreads();
  if (ty == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sh[i][tx];
    out[c] += t;
  }
}

// The same for a table of (in, out, ld, C) jobs sharing one row count: the backward of a step makes ~80 bias gradients
// of this kind, each a 2-6 CTA launch of ~6 us on an otherwise idle GPU.  CTA b serves the job d with
// block_begin[d] <= b < block_begin[d + 1].
__global__ void __launch_bounds__(256)
reduce_rows_multi_kernel(const B2dReduceDesc* __restrict__ descs, int n_descs, int R) {
  __shared__ float sh[8][33];
  int lo = 0, hi = n_descs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (descs[mid].block_begin <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const B2dReduceDesc d = descs[lo];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = ((int)blockIdx.x - d.block_begin) * 32 + tx;
  float a = 0.f;
  if (c < d.C) {
    for (int r0 = ty; r0 < R; r0 += 32) {
      float t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) t[u] = (r0 + 8 * u < R) ? __ldg(d.in + (long long)(r0 + 8 * u) * d.ld + c) : 0.f;
      a += (t[0] + t[1]) + (t[2] + t[3]);
    }
  }
  sh[ty][tx] = a;
  __syncthreads();
  if (ty == 0 && c < d.C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sh[i][tx];
    atomicAdd(d.out + c, t);        // two jobs of one table may share an output (a bias and its tied copy)
  }
}

// ------------------------------------------------------------------ timestep embedding (utils.py:140-158)
// freqs[k] = exp(-ln(max_period) * k / half) is built on the host with the reference's own torch ops
// (bit-exact table); the kernel does the gather, the int->float cast, the product and cos | sin.
__global__ void timestep_embedding_kernel(const int64_t* __restrict__ t, int N, int dim,
                                          const float* __restrict__ freqs, float* __restrict__ out) {
  const int half = dim / 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * dim) return;
  const int n = i / dim, j = i - n * dim;
  float v = 0.f;
  if (j < 2 * half) {
    const int k = j < half ? j : j - half;
    const float a = __fmul_rn((float)t[n], __ldg(freqs + k));
    v = j < half ? cosf(a) : sinf(a);
  }
  out[i] = v;
}

// ------------------------------------------------------------------ gradient scale (fp16 gradient tensors)
// scale2[0] = 2^k with max|x| * 2^k in [target/2, target), scale2[1] = 2^-k.  Chosen on the device: no host sync.
__global__ void __launch_bounds__(256)
absmax_kernel(const float* __restrict__ x, long long n, unsigned int* __restrict__ out_bits) {
  float m = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(x[i]));
  m = warp_max(m);
  __shared__ float sh[8];
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 8) {
    m = sh[threadIdx.x];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffu, m, o));
    if (threadIdx.x == 0 && m == m) atomicMax(out_bits, __float_as_uint(m));   // non-negative floats order as uints
  }
}
// state: [0] overflow events of the previous backward (consumed here), [1] back-off in 1/1024 bits, [2] skipped
// steps, [3] last consumed count -- the device-side form of MixedPrecisionTrainer's lg_loss_scale update
// (fp16_util.py:222-233: -1 after an overflow, +fp16_scale_growth = 1e-3 after a clean step).
__global__ void grad_scale_finalize_kernel(float* __restrict__ scale2, float target, int* __restrict__ state) {
  int backoff_bits = 0;
  if (state != nullptr) {
    const int ov = state[0];
    state[3] = ov;
    if (ov > 0) { state[1] += 1024; state[2] += 1; }
    else if (state[1] > 0) state[1] -= 1;
    state[0] = 0;
    backoff_bits = (state[1] + 1023) >> 10;
  }
  const float m = scale2[0];
  float s = 1.0f;
  if (m > 0.f && m < INFINITY) {
    int e;
    frexpf(target / m, &e);              // target/m = f * 2^e, f in [0.5, 1)  ->  2^(e-1) <= target/m
    e = max(-60, min(60, e - 1 - backoff_bits));
    s = ldexpf(1.0f, e);
  }
  scale2[0] = s;
  scale2[1] = 1.0f / s;
}
__global__ void __launch_bounds__(256)
scale_f32_kernel(float* __restrict__ x, long long n, const float* __restrict__ scale_ptr) {
  const float s = __ldg(scale_ptr);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    x[i] *= s;
}

static unsigned ew_blocks(long long work, int threads = 256) {
  long long b = (work + threads - 1) / threads;
  long long cap = (long long)num_sms() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_assemble_input(const float* xt, const float* cond, const float* wcond, const float* audio, int N,
                                  int S, int Cin, int Cc, int Hc, int Wc, int Cim, int A, int Cpad, void* out,
                                  void* stream) {
  B2D_REQUIRE(xt && out && (A == 0 || audio) && (Cim == 0 || (cond && wcond)), B2D_ERR_SHAPE, "assemble_input: null pointer");
  B2D_REQUIRE(Cpad % 8 == 0 && Cin + Cim + A <= Cpad && Cc <= 8 && aligned16(out), B2D_ERR_SHAPE,
              "assemble_input: channels %d+%d+%d do not fit Cpad=%d", Cin, Cim, A, Cpad);
  AssembleParams p{xt, cond, wcond, audio, N, S, Cin, Cc, Hc, Wc, Cim, A, Cpad};
  B2D_REQUIRE(Cc <= 4 && Cpad / 8 <= 256 && N <= 65535, B2D_ERR_SHAPE, "assemble_input: Cc=%d (max 4), Cpad=%d, N=%d", Cc, Cpad, N);
  long long chunks = ((long long)S * S + 7) / 8;
  const long long cap = ((long long)num_sms() * 32) / (2LL * N);      // ~4 waves of 8 CTAs per SM over both roles
  if (chunks > cap) chunks = cap;
  if (chunks < 1) chunks = 1;
  assemble_input_kernel<<<dim3((unsigned)chunks, (unsigned)N, 2u), 256, 0, (cudaStream_t)stream>>>(p, (__half*)out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_assemble_input_bwd(const void* dX, const float* cond, int N, int S, int Cin, int Cc, int Hc, int Wc,
                                      int Cim, int A, int Cpad, float* d_wcond, float* d_audio, void* stream) {
  B2D_REQUIRE(dX && (Cim == 0 || (cond && d_wcond)) && (A == 0 || d_audio), B2D_ERR_SHAPE, "assemble_input_bwd: null pointer");
  B2D_REQUIRE(Cc <= 3 && N <= 65535, B2D_ERR_SHAPE, "assemble_input_bwd: Cc=%d (max 3), N=%d", Cc, N);
  AssembleParams p{nullptr, cond, nullptr, nullptr, N, S, Cin, Cc, Hc, Wc, Cim, A, Cpad};
  const int V = (Cin + Cim + A + 7) / 8;
  B2D_REQUIRE(V <= 256, B2D_ERR_SHAPE, "assemble_input_bwd: too many channels");
  const int rows = 256 / V;
  long long chunks = ((long long)S * S + rows - 1) / rows;
  long long cap = ((long long)num_sms() * 8) / N;      // 8 CTAs of 256 threads per SM are resident: one wave
  if (chunks > cap) chunks = cap;
  if (chunks < 1) chunks = 1;
  dim3 grid((unsigned)chunks, (unsigned)N);
  assemble_input_bwd_kernel<<<grid, 256, (Cim * Cc + A) * sizeof(float), (cudaStream_t)stream>>>(
      p, (const __half*)dX, d_wcond, d_audio);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_upsample2x(const void* x, void* y, int N, int H, int W, int C, void* stream) {
  B2D_REQUIRE(x && y && C % 8 == 0, B2D_ERR_SHAPE, "upsample2x: bad arguments");
  upsample2x_kernel<<<ew_blocks((long long)N * 4 * H * W * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, (__half*)y, N, H, W, C);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_upsample2x_bwd(const void* dy, void* dx, int N, int H, int W, int C, void* stream) {
  B2D_REQUIRE(dy && dx && C % 8 == 0, B2D_ERR_SHAPE, "upsample2x_bwd: bad arguments");
  upsample2x_bwd_kernel<<<ew_blocks((long long)N * H * W * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)dy, (__half*)dx, N, H, W, C);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_zero_insert2x(const void* dy, void* z, int N, int Ho, int Wo, int C, void* stream) {
  B2D_REQUIRE(dy && z && C % 8 == 0, B2D_ERR_SHAPE, "zero_insert2x: bad arguments");
  zero_insert2x_kernel<<<ew_blocks((long long)N * 4 * Ho * Wo * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)dy, (__half*)z, N, Ho, Wo, C);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_nchw_to_pixel_major(const float* src, void* dst, int N, int C, int HW, int Cpad,
                                       const float* scale, void* stream) {
  B2D_REQUIRE(src && dst && Cpad % 8 == 0 && C <= Cpad, B2D_ERR_SHAPE, "nchw_to_pixel_major: bad arguments");
  nchw_to_pixel_major_kernel<<<ew_blocks((long long)N * HW * (Cpad / 8)), 256, 0, (cudaStream_t)stream>>>(
      src, (__half*)dst, N, C, HW, Cpad, scale);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_absmax(const float* x, int64_t n, float* out1, void* stream) {
  B2D_REQUIRE(x && out1 && n > 0, B2D_ERR_SHAPE, "absmax: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  B2D_CHECK_CUDA(cudaMemsetAsync(out1, 0, sizeof(float), st));
  absmax_kernel<<<ew_blocks(n), 256, 0, st>>>(x, n, (unsigned int*)out1);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_grad_scale_finalize(float* scale2, float target, int32_t* state, void* stream) {
  B2D_REQUIRE(scale2 && target > 0.f, B2D_ERR_SHAPE, "grad_scale_finalize: bad arguments");
  grad_scale_finalize_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(scale2, target, state);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_grad_scale(const float* x, int64_t n, float target, float* scale2, void* stream) {
  B2D_REQUIRE(x && scale2 && n > 0 && target > 0.f, B2D_ERR_SHAPE, "grad_scale: bad arguments");
  int rc = b2d_absmax(x, n, scale2, stream);
  if (rc) return rc;
  return b2d_grad_scale_finalize(scale2, target, nullptr, stream);
}
extern "C" int b2d_scale_f32(float* x, int64_t n, const float* scale, void* stream) {
  B2D_REQUIRE(x && scale && n >= 0, B2D_ERR_SHAPE, "scale_f32: bad arguments");
  if (n == 0) return B2D_OK;
  scale_f32_kernel<<<ew_blocks(n), 256, 0, (cudaStream_t)stream>>>(x, n, scale);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_colsum(const void* x, int64_t ld, int c_off, int C, int N, int HW, float* out, int64_t ldout,
                          void* stream) {
  B2D_REQUIRE(x && out && C % 8 == 0 && C <= 2048 && c_off % 8 == 0 && ld % 8 == 0 && N <= 65535, B2D_ERR_SHAPE,
              "colsum: bad arguments C=%d c_off=%d ld=%lld", C, c_off, (long long)ld);
  const int V = C / 8;
  const int rows = 256 / V > 0 ? 256 / V : 1;
  long long chunks = ((long long)HW + rows - 1) / rows;
  long long cap = ((long long)num_sms() * 8) / N;   // rounded DOWN: 8 CTAs per SM are resident, a few more are a second wave
  if (chunks > cap) chunks = cap;
  if (chunks < 1) chunks = 1;
  dim3 grid((unsigned)chunks, (unsigned)N);
  colsum_kernel<<<grid, V * rows, C * sizeof(float), (cudaStream_t)stream>>>((const __half*)x, ld, c_off, C, HW, out, ldout);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_pack_conv_weight(const float* w, int64_t s_o, int64_t s_i, int64_t s_k, int Cout, int Cin, int K,
                                    int transpose, void* out, int rows_pad, int inner_pad, void* stream) {
  B2D_REQUIRE(w && out && Cout > 0 && Cin > 0 && K > 0, B2D_ERR_SHAPE, "pack_conv_weight: bad arguments");
  B2D_REQUIRE(rows_pad >= (transpose ? Cin : Cout) && inner_pad >= (transpose ? Cout : Cin), B2D_ERR_SHAPE,
              "pack_conv_weight: padded shape (%d, %d) smaller than the weights (%d, %d)", rows_pad, inner_pad, Cout, Cin);
  const long long total = (long long)rows_pad * K * inner_pad;
  pack_conv_weight_kernel<<<ew_blocks(total), 256, 0, (cudaStream_t)stream>>>(w, s_o, s_i, s_k, Cout, Cin, K, transpose,
                                                                                 (__half*)out, rows_pad, inner_pad);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_pack_block_elems(void) { return b2d::PACK_BLOCK_ELEMS; }
extern "C" int b2d_pack_conv_weights_multi(const B2dPackDesc* descs_dev, int n_descs, int total_blocks, void* stream) {
  if (n_descs == 0 || total_blocks == 0) return B2D_OK;
  B2D_REQUIRE(descs_dev && n_descs > 0 && total_blocks > 0, B2D_ERR_SHAPE, "pack_conv_weights_multi: bad arguments");
  B2D_REQUIRE((reinterpret_cast<uintptr_t>(descs_dev) & 7) == 0, B2D_ERR_ALIGN, "pack_conv_weights_multi: table must be 8-byte aligned");
  pack_conv_weights_multi_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>(descs_dev, n_descs);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_reduce_rows(const float* in, int R, int C, int64_t ld, float* out, void* stream) {
  B2D_REQUIRE(in && out && R > 0 && C > 0, B2D_ERR_SHAPE, "reduce_rows: bad arguments");
  reduce_rows_kernel<<<(C + 31) / 32, 256, 0, (cudaStream_t)stream>>>(in, R, C, ld, out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_reduce_rows_multi(const B2dReduceDesc* descs_dev, int n_descs, int R, int total_blocks, void* stream) {
  if (n_descs == 0 || total_blocks == 0) return B2D_OK;
  B2D_REQUIRE(descs_dev && n_descs > 0 && total_blocks > 0 && R > 0, B2D_ERR_SHAPE, "reduce_rows_multi: bad arguments");
  B2D_REQUIRE((reinterpret_cast<uintptr_t>(descs_dev) & 7) == 0, B2D_ERR_ALIGN, "reduce_rows_multi: table must be 8-byte aligned");
  reduce_rows_multi_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>(descs_dev, n_descs, R);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
extern "C" int b2d_timestep_embedding(const int64_t* t, int N, int dim, const float* freqs, float* out, void* stream) {
  B2D_REQUIRE(t && out && freqs && N > 0 && dim > 0, B2D_ERR_SHAPE, "timestep_embedding: bad arguments");
  timestep_embedding_kernel<<<(N * dim + 255) / 256, 256, 0, (cudaStream_t)stream>>>(t, N, dim, freqs, out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
