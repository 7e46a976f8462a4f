// dataset_window.cu -- SURVEY section 8f row N2: the per-item input pipeline of TalkingFaceFrameDataset.__getitem__
// (/root/reference/video-generation/dataset.py:78-136) batched on the device, one CTA per item:
//   frame / audio-window index arithmetic (IEEE double, bit-exact with the reference's Python floats),
//   waveform gather, 300 Hz high-pass biquad + clamp, (x - mean) / std, the reference's 1 Hz -> 16 kHz sinc resample of
//   which only the first 4000 outputs survive, and the feature extractor's zero-mean / unit-variance normalisation;
//   plus the frame gather: uint8 HWC frames 0 and output_frame_idx -> fp32 CHW, ToTensor + Normalize(0.5, 0.5).
// Integer / index work is bit-exact; the float chain follows the reference's operation order in fp32.
#include "common.cuh"

namespace b2d {

constexpr int AW_THREADS = 256;
constexpr int AW_MAX_SEG = 11776;       // samples of one window held in shared memory (46 KB); 5 frames at 8 fps, 16 kHz
constexpr int AW_OUT = 4000;            // dataset.py:125 target_length
constexpr int AW_TAPS = 15;             // 2 * ceil(6 / 0.99) + 1 taps of torchaudio's sinc_interp_hann filter at 1 Hz -> 16 kHz
constexpr int AW_WIDTH = 7;

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < AW_THREADS / 32; ++i) t += red[i];
  return t;
}

// dataset.py:98,112-119 in IEEE double with the reference's operation order (no contraction):
//   idx = min(frame_end, total - 1); d = 1 / fps; start = int(sr * max(0, (idx - 5) * d)); end = int(sr * (idx * d))
__device__ __forceinline__ void window_indices(long long frame_end, long long total_frames, double fps, int sr,
                                               long long* idx, long long* s, long long* e) {
  const long long i = frame_end < total_frames - 1 ? frame_end : total_frames - 1;
  const double d = __ddiv_rn(1.0, fps);
  double start_sec = __dmul_rn((double)(i - 5), d);
  if (!(start_sec > 0.0)) start_sec = 0.0;
  const double end_sec = __dmul_rn((double)i, d);
  *idx = i;
  *s = (long long)__dmul_rn((double)sr, start_sec);      // int(): truncation toward zero
  *e = (long long)__dmul_rn((double)sr, end_sec);
}

__global__ void __launch_bounds__(AW_THREADS)
audio_window_kernel(const float* __restrict__ waveforms, const long long* __restrict__ wave_offset,
                    const long long* __restrict__ wave_len, const int* __restrict__ video_id,
                    const long long* __restrict__ frame_end, const long long* __restrict__ total_frames,
                    const double* __restrict__ fps, const int* __restrict__ sample_rate,
                    const float* __restrict__ sinc_rows, const float* __restrict__ biquad,   // biquad: [n_items][5] b0 b1 b2 a1 a2
                    float* __restrict__ out, long long* __restrict__ index_out, int* __restrict__ err_flag) {
  __shared__ float seg[AW_MAX_SEG];
  __shared__ float red[AW_THREADS / 32];
  __shared__ float zp[AW_TAPS + 1];
  __shared__ long long s_idx[3];
  const int item = blockIdx.x;
  const int vid = video_id[item];
  const long long wlen = wave_len[vid];
  if (threadIdx.x == 0) {
    long long i, s, e;
    window_indices(frame_end[item], total_frames[item], fps[item], sample_rate[item], &i, &s, &e);
    index_out[item * 3 + 0] = i;                           // output_frame_idx (the input frame is always 0: dataset.py:97)
    index_out[item * 3 + 1] = s;
    index_out[item * 3 + 2] = e;
    // Python slice semantics of waveform[:, s:e] (both non-negative here): clip to the length
    if (s > wlen) s = wlen;
    if (e > wlen) e = wlen;
    if (e < s) e = s;
    s_idx[0] = s; s_idx[1] = e;
  }
  __syncthreads();
  const long long s0 = s_idx[0];
  const int L = (int)(s_idx[1] - s0);
  float* o = out + (long long)item * AW_OUT;
  if (L < 2 || L > AW_MAX_SEG) {                           // empty / one-sample window: the reference's std is NaN
    if (threadIdx.x == 0 && err_flag) atomicExch(err_flag, L > AW_MAX_SEG ? 2 : 1);
    for (int i = threadIdx.x; i < AW_OUT; i += AW_THREADS) o[i] = __int_as_float(0x7fc00000);
    return;
  }
  const float* w = waveforms + wave_offset[vid] + s0;
  for (int i = threadIdx.x; i < L; i += AW_THREADS) seg[i] = __ldg(w + i);
  __syncthreads();
  // ---- high-pass biquad (torchaudio lfilter: direct form I in fp32, clamp to [-1, 1] AFTER the recursion): the
  // recursion is sequential; one thread walks the window in shared memory (L <= 11776 steps of ~6 dependent flops).
  if (threadIdx.x == 0) {
    const float b0 = biquad[item * 5 + 0], b1 = biquad[item * 5 + 1], b2 = biquad[item * 5 + 2];
    const float a1 = biquad[item * 5 + 3], a2 = biquad[item * 5 + 4];
    float x1 = 0.f, x2 = 0.f, y1 = 0.f, y2 = 0.f;
    for (int t = 0; t < L; ++t) {
      const float xt = seg[t];
      float acc = __fadd_rn(__fadd_rn(__fmul_rn(b2, x2), __fmul_rn(b1, x1)), __fmul_rn(b0, xt));
      acc = __fsub_rn(acc, __fadd_rn(__fmul_rn(a2, y2), __fmul_rn(a1, y1)));
      seg[t] = fminf(fmaxf(acc, -1.0f), 1.0f);
      x2 = x1; x1 = xt; y2 = y1; y1 = acc;
    }
  }
  __syncthreads();
  // ---- normalize_waveform (dataset.py:34-37): mean and UNBIASED std over the window, two passes
  float a = 0.f;
  for (int i = threadIdx.x; i < L; i += AW_THREADS) a += seg[i];
  const float mean = block_sum(a, red) / (float)L;
  float q = 0.f;
  for (int i = threadIdx.x; i < L; i += AW_THREADS) { const float d = seg[i] - mean; q = fmaf(d, d, q); }
  const float stdv = sqrtf(block_sum(q, red) / (float)(L - 1));
  // ---- process_audio (dataset.py:51-66): Resample(orig_freq = size(0) = 1 channel, new_freq = 16000), then [:4000]:
  // output i < 4000 is filter row i applied to the zero-padded first samples, padded[k] = z[k - 7]
  if (threadIdx.x < AW_TAPS) {
    const int k = threadIdx.x - AW_WIDTH;
    zp[threadIdx.x] = (k >= 0 && k < L) ? (seg[k] - mean) / stdv : 0.f;
  }
  __syncthreads();
  float r[(AW_OUT + AW_THREADS - 1) / AW_THREADS];
  float sa = 0.f;
#pragma unroll
  for (int j = 0; j < (AW_OUT + AW_THREADS - 1) / AW_THREADS; ++j) {
    const int i = threadIdx.x + j * AW_THREADS;
    float acc = 0.f;
    if (i < AW_OUT) {
#pragma unroll
      for (int k = 0; k < AW_TAPS; ++k) acc = fmaf(__ldg(sinc_rows + i * AW_TAPS + k), zp[k], acc);
      sa += acc;
    }
    r[j] = acc;
  }
  // ---- Wav2Vec2FeatureExtractor.zero_mean_unit_var_norm (dataset.py:130): (x - mean) / sqrt(var + 1e-7), population var
  const float m2 = block_sum(sa, red) / (float)AW_OUT;
  float sq = 0.f;
#pragma unroll
  for (int j = 0; j < (AW_OUT + AW_THREADS - 1) / AW_THREADS; ++j) {
    const int i = threadIdx.x + j * AW_THREADS;
    if (i < AW_OUT) { const float d = r[j] - m2; sq = fmaf(d, d, sq); }
  }
  const float inv = 1.0f / sqrtf(block_sum(sq, red) / (float)AW_OUT + 1e-7f);
#pragma unroll
  for (int j = 0; j < (AW_OUT + AW_THREADS - 1) / AW_THREADS; ++j) {
    const int i = threadIdx.x + j * AW_THREADS;
    if (i < AW_OUT) o[i] = (r[j] - m2) * inv;
  }
}

// frames: uint8 (total T of all videos, H, W, 3); out: fp32 (n_items, 2, 3, H, W) = [input frame 0, output frame idx]
// ToTensor: v / 255 ; Normalize(0.5, 0.5): (t - 0.5) / 0.5 -- three separately rounded fp32 operations (bit-exact).
__global__ void __launch_bounds__(256)
frame_gather_kernel(const unsigned char* __restrict__ frames, const long long* __restrict__ video_first_frame,
                    const int* __restrict__ video_id, const long long* __restrict__ index_out, int H, int W,
                    float* __restrict__ out) {
  const int item = blockIdx.y, which = blockIdx.z;
  const long long f = video_first_frame[video_id[item]] + (which == 0 ? 0 : index_out[item * 3]);
  const long long HW = (long long)H * W;
  const unsigned char* src = frames + f * HW * 3;
  float* dst = out + ((long long)item * 2 + which) * 3 * HW;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (long long)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float t = __fdiv_rn((float)src[p * 3 + c], 255.0f);
      dst[c * HW + p] = __fdiv_rn(__fsub_rn(t, 0.5f), 0.5f);
    }
  }
}

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_audio_window_max_samples(void) { return AW_MAX_SEG; }

extern "C" int b2d_audio_window(const float* waveforms, const int64_t* wave_offset, const int64_t* wave_len,
                                const int32_t* video_id, const int64_t* frame_end, const int64_t* total_frames,
                                const double* fps, const int32_t* sample_rate, const float* sinc_rows,
                                const float* biquad, int n_items, float* out, int64_t* index_out, int* err_flag,
                                void* stream) {
  if (n_items == 0) return B2D_OK;
  B2D_REQUIRE(waveforms && wave_offset && wave_len && video_id && frame_end && total_frames && fps && sample_rate &&
              sinc_rows && biquad && out && index_out && n_items > 0, B2D_ERR_SHAPE, "audio_window: null argument");
  audio_window_kernel<<<n_items, AW_THREADS, 0, (cudaStream_t)stream>>>(
      waveforms, (const long long*)wave_offset, (const long long*)wave_len, video_id, (const long long*)frame_end,
      (const long long*)total_frames, fps, sample_rate, sinc_rows, biquad, out, (long long*)index_out, err_flag);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

extern "C" int b2d_frame_gather(const uint8_t* frames, const int64_t* video_first_frame, const int32_t* video_id,
                                const int64_t* index_out, int n_items, int H, int W, float* out, void* stream) {
  if (n_items == 0) return B2D_OK;
  B2D_REQUIRE(frames && video_first_frame && video_id && index_out && out && n_items > 0 && H > 0 && W > 0 &&
              n_items <= 65535, B2D_ERR_SHAPE, "frame_gather: bad arguments");
  const long long HW = (long long)H * W;
  int bx = (int)((HW + 255) / 256);
  if (bx > 64) bx = 64;
  frame_gather_kernel<<<dim3(bx, n_items, 2), 256, 0, (cudaStream_t)stream>>>(
      frames, (const long long*)video_first_frame, video_id, (const long long*)index_out, H, W, out);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}
