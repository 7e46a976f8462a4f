/*
 * b2d.h -- C ABI of libb200diff.so: the sm_100a kernels behind the denoiser
 * train step / sampler step of video-generation/diffusion (reference:
 * wdas03/lipreading-video-generation, paths relative to /root/reference).
 *
 * Contract:
 *  - every call ENQUEUES work on `stream` and returns; the library never
 *    allocates or frees device memory, never keeps a pointer past the call and
 *    never synchronises the stream (CUDA-graph capturable);
 *  - all buffers are caller-owned device pointers (PyTorch caching allocator);
 *  - return 0 on success, a B2D_ERR_* code otherwise; text via b2d_last_error();
 *  - there is NO CPU fallback and no multi-backend dispatch: a device that is
 *    not compute capability 10.x yields B2D_ERR_ARCH.
 *
 * `stream` is a cudaStream_t passed as void* so that the header is plain C.
 */
#ifndef B2D_H_
#define B2D_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2D_OK 0
#define B2D_ERR_SHAPE 1
#define B2D_ERR_ALIGN 2
#define B2D_ERR_ARCH 3
#define B2D_ERR_WORKSPACE 4
#define B2D_ERR_CUDA 5

int b2d_version(void);
const char* b2d_last_error(void);
/* 0 when the current device is sm_100-class, B2D_ERR_ARCH otherwise. */
int b2d_check_device(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
long long b2d_launch_count(void);

/* ------------------------------------------------------------------------
 * Diffusion math (elementwise, HBM bound, fp32).
 * ---------------------------------------------------------------------- */

/* LinearNoiseScheduler.add_noise  (video-generation/diffusion/
 * linear_noise_scheduler.py:24-46):  x_t = sqrt_acp[t[b]]*x0 + sqrt_1macp[t[b]]*eps
 * with one timestep per sample.  x0/eps/xt: (B, per_sample) fp32 contiguous;
 * t: (B,) int64; tables: (T,) fp32 device copies of sqrt_alpha_cum_prod and
 * sqrt_one_minus_alpha_cum_prod.  mul, mul, add are rounded separately
 * (no FMA contraction) so the result is bit-identical to the eager reference.
 * An out-of-range t sets *err_flag (device int32, may be NULL) to 1 and
 * writes NaN for that sample (the reference raises IndexError). */
int b2d_q_sample(const float* x0, const float* eps, const int64_t* t,
                 const float* sqrt_acp, const float* sqrt_1macp, int T,
                 float* xt, int64_t B, int64_t per_sample, int* err_flag,
                 void* stream);

/* Sampler updates.  All three take the per-step scalar coefficients that the
 * host gathers from the schedule tables (bit-exact fp32 table values) so the
 * kernel does the reference's arithmetic in the reference's order.
 * z may be NULL when use_noise == 0.  n = number of elements. */

/* LinearNoiseScheduler.sample_prev_timestep (linear_noise_scheduler.py:48-76)
 *   x0   = clamp((xt - s1m*eps) / sqrt(acp_t), -1, 1)
 *   mean = (xt - beta*eps / s1m) / sqrt(alpha_t)
 *   out  = mean + sigma*z   (t > 0)   |   mean   (t == 0) */
int b2d_ddpm_step(const float* xt, const float* eps, const float* z,
                  float s1m, float sqrt_acp, float beta, float sqrt_alpha,
                  float sigma, int use_noise, float* x_prev, float* x0,
                  int64_t n, void* stream);

/* LinearNoiseSchedulerV2.sample_prev_timestep (linear_noise_scheduler.py:91-101)
 *   mean = xt - (s1m*eps) / sqrt(alpha_t);  out = mean + sigma*z
 *   x0   = clamp((xt - s1m*eps) / sqrt_acp, -1, 1) */
int b2d_ddpm_v2_step(const float* xt, const float* eps, const float* z,
                     float s1m, float sqrt_alpha, float sqrt_acp, float sigma,
                     float* x_prev, float* x0, int64_t n, void* stream);

/* CosineNoiseScheduler.sample_prev_timestep (noise_scheduler.py:13-29)
 *   mean = (xt - s1m*eps) / sqrt_acp;  out = mean + sigma*z (t>0) | mean
 *   second return value is `mean` itself. */
int b2d_cosine_step(const float* xt, const float* eps, const float* z,
                    float s1m, float sqrt_acp, float sigma, int use_noise,
                    float* x_prev, float* mean_out, int64_t n, void* stream);

/* DDIM eta=0 (not in the reference -- parity unpinned; Song et al. eq. 12):
 *   x0 = clamp((xt - s1m_t*eps)/sqrt_acp_t, -1, 1)
 *   out = sqrt_acp_prev*x0 + s1m_prev*eps */
int b2d_ddim_step(const float* xt, const float* eps, float s1m_t,
                  float sqrt_acp_t, float sqrt_acp_prev, float s1m_prev,
                  float* x_prev, float* x0, int64_t n, void* stream);

/* nn.MSELoss() as called at train.py:103,132: mean((pred-target)^2).
 * partial: workspace of b2d_mse_workspace_bytes(n) bytes; loss: 1 fp32.
 * Deterministic two-stage reduction (warp shuffle -> block -> last block). */
size_t b2d_mse_workspace_bytes(int64_t n);
int b2d_mse_fwd(const float* pred, const float* target, int64_t n,
                void* workspace, size_t workspace_bytes, float* loss,
                void* stream);
/* grad_pred = grad_out * 2*(pred-target)/n ; grad_out is a device scalar. */
int b2d_mse_bwd(const float* pred, const float* target, const float* grad_out,
                int64_t n, float* grad_pred, void* stream);

/* ------------------------------------------------------------------------
 * Denoiser kernels.  Activation layout inside the library is PIXEL-MAJOR
 * (N,H,W,C) fp16 -- the channels-last image of the reference's NCHW / (B,C,L)
 * tensors -- so that one row of a tile is one K-major 128-byte line for the
 * tcgen05 tensor cores; accumulation is always fp32.  Gradient tensors are fp16
 * too and carry a power-of-two loss scale (b2d_grad_scale).  (bf16 storage
 * missed north_star's 2e-2 gradient tolerance)
 * Weights stay fp32 OIHW in state_dict; the host side packs fp16 copies
 * [Cout][tap][Cin] for these calls.
 * ---------------------------------------------------------------------- */

/* Implicit-GEMM convolution on tcgen05/TMEM fed by TMA.
 * Replaces nn.Conv2d 3x3 pad 1 (stride 1 or 2) and 1x1 convs / Conv1d k=1 of
 * unet.py:143-152 (Downsample), :194-234 (ResBlock), :298-317 (AttentionBlock
 * qkv / proj_out), :624-628 (head); also used for dgrad with flipped weights.
 *   out[n,ho,wo,co] = sum_{dy,dx,ci} X[n, ho*s+dy-p, wo*s+dx-p, ci] * w[co][dy*k+dx][ci]
 *                     + bias[co] + emb[n,co] + residual[n,ho,wo,co]
 * X is the channel concatenation [x | x2] (torch.cat at unet.py:672) without
 * materialising it.  C1, C2 multiples of 64; W a power of two <= 128. */
typedef struct b2d_conv_args {
  const void* x;   int64_t ldx;  int C1;     /* fp16 (N,H,W,ldx), channels [0,C1) used */
  const void* x2;  int64_t ldx2; int C2;     /* optional second source, or NULL / 0 */
  int N, H, W;                               /* input spatial size */
  int ksize;                                 /* 1 or 3 (padding ksize/2) */
  int stride;                                /* 1 or 2 */
  const void* w;                             /* fp16 [CoutPad][ksize*ksize*(C1+C2)] */
  int Cout;                                  /* logical output channels */
  int CoutPad;                               /* rows of w: multiple of 16 */
  const float* bias;                         /* [Cout] or NULL */
  const float* emb; int64_t ldemb;           /* [N][ldemb] per-sample add (unet.py:266) or NULL */
  const void* residual; int64_t ldres;       /* fp16 (N,Ho,Wo,ldres) added in the epilogue, or NULL */
  void* out; int64_t ldo;                    /* fp16 (N,Ho,Wo,ldo) or NULL */
  float* out_nchw_f32;                       /* optional fp32 (N,Cout,Ho,Wo) output (model output) */
  /* ---- optional GroupNorm pieces fused into the epilogue (north_star item 2: "fused GroupNorm/SiLU/residual
   * epilogues"; utils.py:54-56, unet.py:194-198,218-225,297).  Both need the staged epilogue (Cout % 64 == 0, fp16
   * `out`), Ho*Wo % 32 == 0 and (images per 128-pixel tile) * (channels per tile) <= 256; b2d_conv2d_gn_fusable()
   * tells.  All NULL / 0 = plain convolution. */
  /* (1) statistics of the OUTPUT for the GroupNorm that consumes it: gn_stats_out[n][c/2][0..1] += sum / sum of
   * squares over the pixels of image n of output channels (c, c+1), from the fp32 epilogue values.  fp64
   * [N][Cout/2][2], caller zeroes.  The consumer (b2d_groupnorm_apply) forms its 32 group moments from these pairs,
   * also across the two tensors of a skip concatenation -- the separate statistics pass over the tensor disappears. */
  double* gn_stats_out;
  /* (2) this convolution is the DATA GRADIENT that produces dL/dy for y = act(GroupNorm(x)) [* dropout mask]: the
   * epilogue turns the accumulator dy into dyh = dy * act'(z) * mask/(1-p) (z = xhat*gamma + beta, from the saved mean
   * / rstd and x loaded by TMA exactly like a residual tile), stores dyh in `out`, and accumulates
   * gnb_ab[n][c][0..1] += (sum_pixels dyh, sum_pixels dyh * x): the reduce pass of the GroupNorm backward.
   * x = [gnb_x1 | gnb_x2] (C1 multiple of 64, C1 + C2 == Cout); mean / rstd fp32 [N][32]; gamma / beta fp32 [Cout];
   * gnb_mask: dropout keep bits written by b2d_groupnorm_apply ((N, Ho*Wo, Cout/8) bytes) or NULL. */
  const void* gnb_x1; int64_t gnb_ld1; int gnb_C1;
  const void* gnb_x2; int64_t gnb_ld2;
  const float* gnb_mean; const float* gnb_rstd;
  const float* gnb_gamma; const float* gnb_beta;
  int gnb_act;
  const uint8_t* gnb_mask; float gnb_keep_scale;
  float* gnb_ab;                             /* fp32 [N][Cout][2], accumulated (caller zeroes) */
} b2d_conv_args;

int b2d_conv2d(const b2d_conv_args* a, void* stream);
/* 1 when a convolution with this output plane / channel count can carry the fused GroupNorm pieces above. */
int b2d_conv2d_gn_fusable(int Ho, int Wo, int Cout);

/* Weight gradient of the same convolution (fp32, ACCUMULATED into dw):
 *   dw[co][tap][ci] += sum_{n,ho,wo} dy[n,ho,wo,co] * X[n, ho*s+dy-p, wo*s+dx-p, ci]
 * dy: fp16 (N,Ho,Wo,lddy).  dw: fp32 [Cout][taps*(C1+C2)] (caller zeroes). */
typedef struct b2d_wgrad_args {
  const void* x;   int64_t ldx;  int C1;
  const void* x2;  int64_t ldx2; int C2;
  int N, H, W;
  int ksize, stride;
  const void* dy;  int64_t lddy; int Cout;   /* Cout multiple of 64 (or <= 16 via the small path) */
  float* dw;
} b2d_wgrad_args;

int b2d_conv2d_wgrad(const b2d_wgrad_args* a, void* stream);

/* Flash-style self-attention forward (QKVAttentionLegacy.forward unet.py:349-366 when
 * new_order == 0; QKVAttention.forward unet.py:382-401 when new_order == 1):
 *   a = softmax_fp32((q*ch^-1/4)^T (k*ch^-1/4)) v   per (sample, head), L = H*W tokens.
 * qkv: fp16 (N, L, ldqkv) pixel-major image of the reference's (N, 3*heads*ch, L) tensor, with the
 * reference's channel order ([head][q|k|v][ch] legacy, [q|k|v][head][ch] new).
 * out: fp16 (N, L, ldo), head h at channels [h*ch, (h+1)*ch).  lse: fp32 (N, heads, L) or NULL,
 * log2-domain log-sum-exp of the scaled logits (saved for the backward).  ch in {64,128,256}. */
int b2d_attention_fwd(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order,
                      void* out, int64_t ldo, float* lse, void* stream);
/* The same for small grids (the reference samples ONE frame at a time, test.py:54: a 128x128 frame is 64 CTAs of 256
 * queries on 148 SMs): the key range is split over `splits` (1..8) CTAs per query block, each writes its normalised fp32
 * output and log-sum-exp to ws (b2d_attention_fwd_splitkv_ws_floats floats, 16-byte aligned), a second kernel merges
 * them.  ch in {64, 128}, L > 128.  Equal to b2d_attention_fwd up to the rounding of the 16-bit output. */
int64_t b2d_attention_fwd_splitkv_ws_floats(int N, int L, int heads, int ch, int splits);
int b2d_attention_fwd_splitkv(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order,
                              int splits, float* ws, int64_t ws_floats, void* out, int64_t ldo, float* lse,
                              void* stream);

/* GroupNorm32(32, C) [+ SiLU] [+ dropout] on pixel-major fp16 (utils.py:54-56,130-137; call sites
 * unet.py:194-198 in_layers, :218-225 out_layers incl. nn.Dropout, :297 attention norm, :624-628 head).
 * Statistics in fp32/fp64 as the reference's `.float()` GroupNorm.  The input is the channel
 * concatenation [x1 | x2] (x2 may be NULL).  act: 0 none, 1 SiLU.  drop_p > 0 applies an inverted
 * dropout mask from a counter-based Philox stream keyed by `seed` (re-generated in the backward). */
typedef struct b2d_gn_args {
  const void* x1; int64_t ld1; int C1;
  const void* x2; int64_t ld2; int C2;
  int N, HW;
  const float* gamma; const float* beta;     /* [C1+C2] */
  float eps;
  int act;
  float drop_p;
  uint64_t seed;
  /* backward calls only (may be NULL): int32 device counter, += 1 per thread that saw a gradient element with
   * |value| >= 65504 (a saturated 16-bit gradient) or NaN -- the overflow check of the reference's low-precision
   * trainer (fp16_util.py:236-237 check_overflow), kept on the device: b2d_grad_scale_finalize / b2d_adam_step_guarded
   * consume it without a host synchronisation. */
  int32_t* overflow;
} b2d_gn_args;

size_t b2d_groupnorm_workspace_bytes(int N, int Ctot);
/* y: fp16 (N,HW,C1+C2) contiguous; mean, rstd: fp32 [N][32] (saved for the backward). */
int b2d_groupnorm_fwd(const b2d_gn_args* a, void* y, float* mean, float* rstd, void* workspace,
                      size_t workspace_bytes, void* stream);
/* b2d_groupnorm_fwd without a caller workspace and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (under CUDA graph capture, tight serving
 * loops).  Same arguments, results and errors as b2d_groupnorm_fwd; the N * 64 doubles of scratch live in a
 * file-scope `static __device__` array of B2D_GN_STATIC_MAX_N * 512 bytes (512 KiB per GPU that loads the library,
 * held until exit), whose device address is taken with cudaGetSymbolAddress once per device.  N > B2D_GN_STATIC_MAX_N
 * returns B2D_ERR_WORKSPACE.  Calls no cudaMalloc / cudaFree / allocating routine.  Every call shares the one array
 * (cleared on `stream` before use): calls that overlap on the GPU -- different streams without an ordering between
 * them -- race on it. */
#define B2D_GN_STATIC_MAX_N 1024
int b2d_groupnorm_fwd_static(const b2d_gn_args* a, void* y, float* mean, float* rstd, void* stream);
/* dy: fp16 (N,HW,C1+C2) gradient w.r.t. y.  dx1 (N,HW,lddx1) / dx2 (N,HW,lddx2) receive the input
 * gradients; add_a / add_b (optional, fp16, channel pitch lda / ldb) are ADDED to dx1 so that the
 * residual / skip gradient accumulation is fused; with add_a_full != 0, add_a spans all C1+C2
 * channels and its upper C2 channels are added to dx2.  dgamma/dbeta: fp32 [C1+C2], accumulated.
 * dx1_colsum (optional, fp32 [N][ld_colsum], ACCUMULATED): per-sample sums over pixels of dx1's C1 channels -- the
 * bias / timestep-embedding gradient of the convolution whose output gradient dx1 is (unet.py:266), produced while
 * dx1 is still in registers instead of by a second pass over it (b2d_colsum). */
int b2d_groupnorm_bwd(const b2d_gn_args* a, const void* dy, const float* mean, const float* rstd,
                      void* dx1, int64_t lddx1, void* dx2, int64_t lddx2, const void* add_a, int64_t lda,
                      const void* add_b, int64_t ldb, int add_a_full, float* dx1_colsum, int64_t ld_colsum,
                      float* dgamma, float* dbeta, void* workspace, size_t workspace_bytes, void* stream);

/* ---- GroupNorm with producer-fused statistics (the path the engine uses; the two calls above remain for inputs
 * that no fused convolution produced).
 * b2d_pair_stats: st[n][c/2][0..1] += (sum, sum of squares) over the HW pixels of sample n of channels (c, c+1) of
 *   x fp16 (N, HW, ld): the same quantity b2d_conv2d's gn_stats_out emits (fp64 [N][C/2][2], caller zeroes).
 * b2d_groupnorm_apply: y = [dropout](act(GroupNorm([x1 | x2]))) in ONE pass over the input: the group moments come
 *   from the pair sums st1 (for x1) / st2 (for x2, NULL when there is no x2).  rows_per_sample > 1: sample n of this
 *   call is rows_per_sample consecutive rows of the statistics (dims=3: the T frames of a clip were convolved as
 *   separate images).  mean / rstd fp32 [N][32] are written for the backward.  mask_out (optional, (N, HW, Ctot/8)
 *   bytes): dropout keep bits, bit i of byte j = channel 8j + i, for the fused backward.
 * b2d_groupnorm_bwd_apply: the input-gradient pass given dyh = dy * act'(z) * mask (what the fused data-gradient
 *   epilogue stores) and ab[n][c] = (sum dyh, sum dyh * x) (fp32 [N][Ctot][2]):
 *     dx = rstd * (dyh*gamma - mean_g(dyh*gamma) - xhat * mean_g(dyh*gamma*xhat))  (+ add_a + add_b as above),
 *   and dgamma[c] += sum_n rstd*(ab1 - mean*ab0), dbeta[c] += sum_n ab0.  a->overflow counts saturated gradients. */
int b2d_pair_stats(const void* x, int64_t ld, int C, int N, int HW, double* st, void* stream);
int b2d_groupnorm_apply(const b2d_gn_args* a, const double* st1, const double* st2, int rows_per_sample, void* y,
                        float* mean, float* rstd, uint8_t* mask_out, void* stream);
int b2d_groupnorm_bwd_apply(const b2d_gn_args* a, const void* dyh, const float* mean, const float* rstd,
                            const float* ab, void* dx1, int64_t lddx1, void* dx2, int64_t lddx2, const void* add_a,
                            int64_t lda, const void* add_b, int64_t ldb, int add_a_full, float* dx1_colsum,
                            int64_t ld_colsum, float* dgamma, float* dbeta, void* stream);

/* ------------------------------------------------------------------------
 * Layout / conditioning helpers (HBM bound).
 * ---------------------------------------------------------------------- */

/* UNetAudio.forward conditioning assembly (unet_audio.py:53-61) fused with the NCHW->pixel-major
 * conversion:  out[n,h,w,:] = [ x_t (Cin) | cond_conv_in(nearest(cond)) (Cim) | audio[n] (A) | 0 ... ]
 * xt fp32 (N,Cin,S,S); cond fp32 (N,Cc,Hc,Wc) resized with F.interpolate's default 'nearest';
 * wcond fp32 (Cim,Cc) = cond_conv_in.weight; audio fp32 (N,A); out fp16 (N,S,S,Cpad). */
int b2d_assemble_input(const float* xt, const float* cond, const float* wcond, const float* audio, int N,
                       int S, int Cin, int Cc, int Hc, int Wc, int Cim, int A, int Cpad, void* out,
                       void* stream);
/* Adjoint: d_wcond (Cim,Cc) and d_audio (N,A) are ACCUMULATED from dX fp16 (N,S,S,Cpad). */
int b2d_assemble_input_bwd(const void* dX, const float* cond, int N, int S, int Cin, int Cc, int Hc, int Wc,
                           int Cim, int A, int Cpad, float* d_wcond, float* d_audio, void* stream);

/* Upsample.forward's F.interpolate(scale_factor=2, mode="nearest") (unet.py:112-122) and its adjoint. */
int b2d_upsample2x(const void* x, void* y, int N, int H, int W, int C, void* stream);
int b2d_upsample2x_bwd(const void* dy, void* dx, int N, int H, int W, int C, void* stream);
/* z[n,2h,2w,:] = dy[n,h,w,:], zero elsewhere (stride-2 Downsample dgrad as a stride-1 conv). */
int b2d_zero_insert2x(const void* dy, void* z, int N, int Ho, int Wo, int C, void* stream);
/* fp32 (N,C,HW) -> fp16 (N,HW,Cpad) with zero-padded channels, multiplied by *scale (device scalar) when
 * scale != NULL (the loss scale of the incoming gradient d loss / d pred). */
int b2d_nchw_to_pixel_major(const float* src, void* dst, int N, int C, int HW, int Cpad, const float* scale,
                            void* stream);
/* Loss scale for the fp16 gradient tensors, chosen on the device (no host sync): scale2[0] = 2^k such that
 * max|x| * 2^k lies in [target/2, target), scale2[1] = 2^-k.  (The reference's own low-precision design is fp16
 * with dynamic loss scaling: fp16_util.py:148-233 MixedPrecisionTrainer.) */
int b2d_grad_scale(const float* x, int64_t n, float target, float* scale2, void* stream);
/* The same in two halves, so that ranks which must agree on the scale (one clip split across GPUs: their partial
 * dK / dV are summed in fp16) can all-reduce(MAX) the maximum in between, and with the overflow state of
 * fp16_util.py:148-233 (MixedPrecisionTrainer: lg_loss_scale -= 1 after an overflowed step, += 1e-3 after a clean one):
 *   b2d_absmax:  out1[0] = max |x|   (fp32 bits; zeroed first)
 *   b2d_grad_scale_finalize: scale2[0] holds max |x| on entry; on exit scale2 = (2^k, 2^-k) with k lowered by
 *     ceil(backoff) bits.  state (int32[4], may be NULL): [0] overflow events counted during the PREVIOUS backward
 *     (b2d_gn_args.overflow points here) -- consumed: > 0 raises the back-off by one bit and counts a skipped step,
 *     0 lowers it by 1/1024 bit; [1] back-off in 1/1024 bits; [2] steps skipped so far; [3] copy of [0] for logging. */
int b2d_absmax(const float* x, int64_t n, float* out1, void* stream);
int b2d_grad_scale_finalize(float* scale2, float target, int32_t* state, void* stream);
/* x[i] *= *scale  (unscale the fp32 parameter gradients). */
int b2d_scale_f32(float* x, int64_t n, const float* scale, void* stream);
/* fp32 convolution weights (state_dict layout, element strides s_o / s_i / s_k over Cout / Cin / the K = kh*kw
 * taps; a Conv3d temporal slice is a strided view) -> the fp16 operand layout of b2d_conv2d:
 *   transpose == 0 (fprop):  out[co][k][ci]      = w[co][ci][k]        rows_pad >= Cout, inner_pad >= Cin
 *   transpose == 1 (dgrad):  out[ci][K-1-k][co]  = w[co][ci][k]        rows_pad >= Cin,  inner_pad >= Cout
 * zero padded, saturated to +-65504.  Replaces the host side's permute / clamp / cast / scatter chain
 * (~10 PyTorch launches per convolution per optimiser step). */
int b2d_pack_conv_weight(const float* w, int64_t s_o, int64_t s_i, int64_t s_k, int Cout, int Cin, int K,
                         int transpose, void* out, int rows_pad, int inner_pad, void* stream);
/* The same for EVERY convolution of the model in one launch (the refresh of the operand copies after an optimiser
 * step): descs_dev is a device array of n_descs descriptors; descriptor d owns the CTAs
 * [block_begin[d], block_begin[d+1]) with ceil(rows_pad*K*inner_pad / b2d_pack_block_elems()) CTAs each, block_begin
 * ascending from 0; total_blocks = their sum.  Output buffers are written in place (stable addresses). */
typedef struct B2dPackDesc {
  const float* w;
  void* out;
  int64_t s_o, s_i, s_k;
  int32_t Cout, Cin, K, transpose, rows_pad, inner_pad;
  int32_t block_begin, reserved;
} B2dPackDesc;
int b2d_pack_block_elems(void);
int b2d_pack_conv_weights_multi(const B2dPackDesc* descs_dev, int n_descs, int total_blocks, void* stream);
/* Audio encoder front end.  unet_audio.py:10-18 instantiates HF transformers' Wav2Vec2Model (third-party, not in
 * the reference tree); the first feature-extractor layer (`Wav2Vec2GroupNormConvLayer`: Conv1d(1 -> C, kernel K = 10,
 * stride, no padding) -> GroupNorm(C, C) -> erf GELU) produces the encoder's largest activation, (N, C, T) with
 * T = (S - K) / stride + 1 <= 800.  These two calls replace the layer's conv / norm / activation library ops:
 *   fwd:  out (N, C, T) in out_dtype (0 = fp32, 1 = bf16, 2 = fp16), mean / rstd (N, C) fp32 kept for the backward;
 *   bwd:  dw (C, K), dbias (C, may be NULL), dgamma (C), dbeta (C) are ACCUMULATED (+=); the waveform gets no gradient.
 * audio (N, S) fp32 with row stride ld_audio; bias may be NULL (wav2vec2-base has conv_bias = false). */
int b2d_w2v_layer0_fwd(const float* audio, int64_t ld_audio, const float* w, const float* bias, const float* gamma,
                       const float* beta, int N, int S, int C, int K, int stride, float eps, void* out, int out_dtype,
                       float* mean, float* rstd, void* stream);
int b2d_w2v_layer0_bwd(const float* audio, int64_t ld_audio, const float* w, const float* bias, const float* gamma,
                       const float* beta, int N, int S, int C, int K, int stride, float eps, const void* dout,
                       int dout_dtype, const float* mean, const float* rstd, float* dw, float* dbias, float* dgamma,
                       float* dbeta, void* stream);

/* nn.LayerNorm over the last dimension as HF wav2vec2 calls it 26 times per forward (`Wav2Vec2EncoderLayer.layer_norm`
 * / `.final_layer_norm`, `Wav2Vec2Encoder.layer_norm`, `Wav2Vec2FeatureProjection.layer_norm`; third-party, instantiated
 * at unet_audio.py:14).  x: (R, C) contiguous, x_dtype 0 = fp32, 1 = bf16, 2 = fp16; C / 32 in {2, 4, 8, 16, 24, 32}.
 *   fwd: y fp32 (R, C) = (x - mean) * rstd * gamma + beta; mean / rstd fp32 (R,) kept for the backward;
 *   bwd: dx fp32 (R, C) written; dgamma / dbeta fp32 (C,) ACCUMULATED (+=). */
int b2d_layernorm_fwd(const void* x, int x_dtype, const float* gamma, const float* beta, int R, int C, float eps,
                      float* y, float* mean, float* rstd, void* stream);
int b2d_layernorm_bwd(const void* x, int x_dtype, const float* gamma, const float* mean, const float* rstd,
                      const float* dy, int R, int C, float* dx, float* dgamma, float* dbeta, void* stream);
/* Bias gradient of an nn.Linear of the same encoder: out[c] = sum_r x[r][c] for a 16-bit (R, C) gradient with row
 * stride ld (elements); x_dtype 1 = bf16, 2 = fp16; out fp32 (C,) is OVERWRITTEN (zeroed on the stream first). */
int b2d_colsum_rows16(const void* x, int x_dtype, int64_t ld, int R, int C, float* out, void* stream);
/* out (B, C, R) = in (B, R, C) transposed, 16-bit elements (bf16 / fp16 bit patterns), contiguous, 4-byte aligned: the
 * layout change between the fused layer 0 ((N, C, T) rows) and the channels-last tensor of feature layers 1..6. */
int b2d_transpose16(const void* in, void* out, int B, int R, int C, void* stream);
/* Self-attention over very short sequences (L <= 16 tokens, head dim D = 64): the attention function behind HF
 * transformers' `Wav2Vec2Attention` (the audio encoder of unet_audio.py:10-18 sees 12 frames per 4000-sample window).
 * q / k / v: (B, H, L, D) views with element strides (sb, sh, sl, 1); out, dout, dq, dk, dv: contiguous (B, L, H, D);
 * dtype 0 = fp32, 1 = bf16, 2 = fp16 for all tensors.  Dropout with probability drop_p acts on the probabilities
 * (Philox keyed by `seed`; the backward must be given the forward's seed: it recomputes the mask). */
int b2d_short_attention_fwd(const void* q, const void* k, const void* v, int64_t sb, int64_t sl, int64_t sh, int B,
                            int H, int L, int D, int dtype, float scale, float drop_p, uint64_t seed, void* out,
                            void* stream);
int b2d_short_attention_bwd(const void* q, const void* k, const void* v, int64_t sb, int64_t sl, int64_t sh,
                            const void* dout, int B, int H, int L, int D, int dtype, float scale, float drop_p,
                            uint64_t seed, void* dq, void* dk, void* dv, void* stream);
/* out[n*ldout + c] += sum_pix x[n,pix,c_off+c]  (x fp16 (N,HW,ld)): emb_out / bias / audio gradients. */
int b2d_colsum(const void* x, int64_t ld, int c_off, int C, int N, int HW, float* out, int64_t ldout, void* stream);
/* out[c] += sum_r in[r*ld + c] */
int b2d_reduce_rows(const float* in, int R, int C, int64_t ld, float* out, void* stream);
/* The same for n_descs jobs with one row count R in ONE launch (the ~80 bias gradients of a backward pass:
 * out[c] += sum_r in[r][c] each).  descs_dev: device table; job d owns CTAs [block_begin[d], block_begin[d+1]) with
 * ceil(C / 32) CTAs, block_begin ascending from 0, total_blocks = their sum.  Outputs are accumulated atomically, so two
 * jobs may share an output. */
typedef struct B2dReduceDesc {
  const float* in;
  float* out;
  int64_t ld;
  int32_t C, block_begin;
} B2dReduceDesc;
int b2d_reduce_rows_multi(const B2dReduceDesc* descs_dev, int n_descs, int R, int total_blocks, void* stream);
/* timestep_embedding (utils.py:140-158): out[n] = [cos(t*f) | sin(t*f)], f = host-built table [dim/2]. */
int b2d_timestep_embedding(const int64_t* t, int N, int dim, const float* freqs, float* out, void* stream);

/* Small fp32 Linear layers (time_embed unet.py:483-487, emb_layers unet.py:211-217,
 * AudioFeatureTransformer unet_audio.py:21-30):
 *   y = act_out( f_in(x) @ w^T + bias ), f_in = SiLU if silu_in, act_out = ReLU if relu_out. */
int b2d_linear_fwd(const float* x, const float* w, const float* bias, float* y, int M, int K, int Nout,
                   int silu_in, int relu_out, void* stream);
/* dx written (may be NULL); dw, db ACCUMULATED (db may be NULL). y needed only when relu_out. */
int b2d_linear_bwd(const float* x, const float* w, const float* y, const float* dy, float* dx, float* dw,
                   float* db, int M, int K, int Nout, int silu_in, int relu_out, void* stream);

/* Backward of b2d_attention_fwd w.r.t. the packed qkv tensor (replaces autograd through
 * unet.py:356-366 and the checkpoint recompute of utils.py:179-207).  out / lse are the forward's
 * results; d_out: fp16 (N,L,lddo) gradient w.r.t. out; dqkv: fp16 (N,L,lddqkv) fully written;
 * delta_ws: fp32 workspace of b2d_attention_bwd_ws_floats(N, L, heads, ch) elements (the per-row delta, plus
 * the fp32 dQ accumulator of the single-pass kernel where that one is used). */
int64_t b2d_attention_bwd_ws_floats(int N, int L, int heads, int ch);
int b2d_attention_bwd(const void* qkv, int64_t ldqkv, const void* out, int64_t ldo, const void* d_out,
                      int64_t lddo, const float* lse, int N, int L, int heads, int ch, int new_order,
                      void* dqkv, int64_t lddqkv, float* delta_ws, void* stream);

/* Sequence-parallel shards of the two calls above (one clip's attention over T*H*W positions
 * split across GPUs by QUERY rows; every rank holds the full qkv, so no K/V exchange is needed).
 *   fwd_rows: only query rows [q_row0, q_row0 + q_rows) are computed against all L keys (256-aligned, head dim 64 / 128,
 *             L > 128); out / lse rows outside the range are left untouched -- an all-gather of `out` completes it.
 *   bwd_rows: 128-aligned rows; head dim 64 (L > 128: the fused single-pass kernel) or 128 / 256 (the two-kernel
 *             variant).  dqkv is fully written: Q columns = dQ for the shard's rows and zero elsewhere, K / V columns =
 *             the shard's PARTIAL dK / dV; the sum of the shards' dqkv tensors (one all-reduce) is the full gradient.
 *             `out`, `d_out`, `lse` are only read on the shard's rows (delta is evaluated for every row, the other
 *             rows' values are never used). */
int b2d_attention_fwd_rows(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order,
                           int q_row0, int q_rows, void* out, int64_t ldo, float* lse, void* stream);
int b2d_attention_bwd_rows(const void* qkv, int64_t ldqkv, const void* out, int64_t ldo, const void* d_out,
                           int64_t lddo, const float* lse, int N, int L, int heads, int ch, int new_order,
                           int q_row0, int q_rows, void* dqkv, int64_t lddqkv, float* delta_ws, void* stream);

/* ------------------------------------------------------------------------
 * Optimiser step.
 * ---------------------------------------------------------------------- */

/* One run of consecutive fp32 elements of one parameter tensor and its gradient / Adam moments. */
typedef struct B2dAdamChunk {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  int64_t n;
} B2dAdamChunk;

/* `optimizer.step()` of `torch.optim.Adam(model.parameters(), 1e-2)` (video-generation/diffusion/train.py:102,134)
 * for every parameter tensor in ONE launch: chunks_dev is a device array of n_chunks descriptors (one CTA each; the
 * caller splits tensors into runs of <= 65536 elements).  Update, in torch's evaluation order
 * (torch/optim/adam.py _single_tensor_adam, amsgrad = maximize = False):
 *   g += weight_decay * p;  m = m + (1-beta1)(g - m);  v = beta2 v + (1-beta2) g^2;
 *   p -= (lr / bias_correction1) * m / (sqrt(v) / sqrt(bias_correction2) + eps)
 * with bias_correction{1,2} = 1 - beta{1,2}^step computed by the caller in double (1 - beta is formed in double and
 * rounded once, as torch does: 1.f - 0.999f is 4.7e-5 off).  All arrays fp32, updated in place. */
int b2d_adam_step(const B2dAdamChunk* chunks_dev, int n_chunks, double lr, double beta1, double beta2, float eps,
                  float weight_decay, double bias_correction1, double bias_correction2, void* stream);
/* The same, but a no-op when *found_overflow != 0 (int32 device word, e.g. state[0] above): the skipped step of
 * fp16_util.py:222-233 (`optimize_fp16`: "Found NaN, decreased lg_loss_scale") decided on the device. */
int b2d_adam_step_guarded(const B2dAdamChunk* chunks_dev, int n_chunks, double lr, double beta1, double beta2, float eps,
                          float weight_decay, double bias_correction1, double bias_correction2,
                          const int32_t* found_overflow, void* stream);

/* ------------------------------------------------------------------------
 * Input pipeline: TalkingFaceFrameDataset.__getitem__,
 * video-generation/dataset.py:78-136, batched, one CTA per item.
 * ---------------------------------------------------------------------- */

/* Audio window of every item (dataset.py:98,112-130).  Decoded mono tracks of all videos are concatenated in
 * `waveforms` (fp32), video v occupying [wave_offset[v], wave_offset[v] + wave_len[v]).  Per item i: video_id[i],
 * frame_end[i], total_frames[i], fps[i] (double), sample_rate[i].  Index arithmetic in IEEE double, bit-exact with
 * the reference's Python floats:
 *   idx = min(frame_end, total_frames - 1);  d = 1 / fps;
 *   start = int(sr * max(0, (idx - 5) * d));  end = int(sr * (idx * d))        -> index_out[i] = (idx, start, end)
 * then waveform[start:end] (clipped like a Python slice) -> high-pass biquad with biquad[i] = (b0, b1, b2, a1, a2) / a0
 * (torchaudio.functional.highpass_biquad at 300 Hz, dataset.py:30-32; clamp to [-1, 1] after the recursion) ->
 * (x - mean) / unbiased std (dataset.py:34-37) -> the reference's Resample(orig_freq = 1, new_freq = 16000) of which
 * dataset.py:62-64 keeps outputs [0, 4000): out = sinc_rows (4000 x 15 fp32, filters 0..3999 of torchaudio's
 * sinc_interp_hann kernel) applied to the zero-padded first 8 samples -> (x - mean) / sqrt(var + 1e-7)
 * (Wav2Vec2FeatureExtractor, dataset.py:130).  out: fp32 (n_items, 4000).  A window shorter than 2 samples sets
 * *err_flag = 1 and writes NaN (the reference's std is NaN there); one longer than b2d_audio_window_max_samples()
 * sets 2. */
int b2d_audio_window_max_samples(void);
int b2d_audio_window(const float* waveforms, const int64_t* wave_offset, const int64_t* wave_len,
                     const int32_t* video_id, const int64_t* frame_end, const int64_t* total_frames,
                     const double* fps, const int32_t* sample_rate, const float* sinc_rows, const float* biquad,
                     int n_items, float* out, int64_t* index_out, int* err_flag, void* stream);
/* Frame gather + transforms (dataset.py:97-110, diffusion/train.py:70-75 without the host-side Resize): frames uint8
 * (all videos concatenated, H, W, 3); for item i the input frame is frame 0 of its video and the output frame is
 * index_out[i][0]; out fp32 (n_items, 2, 3, H, W) = ((v / 255) - 0.5) / 0.5, each operation rounded separately
 * (bit-exact with ToTensor + Normalize). */
int b2d_frame_gather(const uint8_t* frames, const int64_t* video_first_frame, const int32_t* video_id,
                     const int64_t* index_out, int n_items, int H, int W, float* out, void* stream);

/* B2D_DENOISER_ABI_BEGIN -- further denoiser kernels are declared below */

#ifdef __cplusplus
}
#endif
#endif /* B2D_H_ */
