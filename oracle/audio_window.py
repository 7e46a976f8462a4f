"""TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the per-item input pipeline of
`TalkingFaceFrameDataset.__getitem__` (/root/reference/video-generation/dataset.py:78-136) -- SURVEY section 8f row N2.

`dataset.py` itself cannot be imported here (decord / moviepy are absent, and it loads a HF processor from the network
at import time), so the pinning is against the reference's CALL SITES evaluated with the reference's own third-party
functions, which are installed: `torchaudio.functional.highpass_biquad` (dataset.py:30-32),
`torchaudio.transforms.Resample` (dataset.py:53-55), `Wav2Vec2FeatureExtractor` normalisation (dataset.py:130, the
feature-extractor half of `Wav2Vec2Processor`; wav2vec2-base-960h's preprocessor_config is feature_size 1, 16 kHz,
do_normalize true, padding_value 0) and `torchvision.transforms` ToTensor / Normalize (train.py:70-75).
`reference_item()` below is that literal restatement (used by oracle/make_golden_audio.py to produce
tests/golden/audio_window.pt); the plain functions are the algorithm restated without those libraries.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module."""
from __future__ import annotations

import math

import numpy as np
import torch

BUFFER_FRAMES = 5          # dataset.py:28
TARGET_LENGTH = 4000       # dataset.py:125
TARGET_SR = 16000          # dataset.py:126


def window_indices(frame_end: int, total_frames: int, video_fps: float, sr: int):
    """dataset.py:98,112-119 -- Python float (IEEE double) arithmetic, truncation by int()."""
    output_frame_idx = min(frame_end, total_frames - 1)
    frame_duration = 1 / video_fps
    start_sec = max(0, (output_frame_idx - BUFFER_FRAMES) * frame_duration)
    end_sec = output_frame_idx * frame_duration
    return output_frame_idx, int(sr * start_sec), int(sr * end_sec)


def highpass_coeffs(sr: int, cutoff: float = 300.0, Q: float = 0.707):
    """torchaudio.functional.highpass_biquad's coefficients (dataset.py:30-32), normalised by a0, as fp32."""
    w0 = 2 * math.pi * cutoff / sr
    alpha = math.sin(w0) / 2.0 / Q
    b0 = (1 + math.cos(w0)) / 2
    b1 = -1 - math.cos(w0)
    b2 = b0
    a0 = 1 + alpha
    a1 = -2 * math.cos(w0)
    a2 = 1 - alpha
    t = torch.tensor([b0, b1, b2, a0, a1, a2], dtype=torch.float32)     # torchaudio builds fp32 tensors, then divides
    b = t[:3] / t[3]
    a = t[3:] / t[3]
    return [float(b[0]), float(b[1]), float(b[2])], [float(a[1]), float(a[2])]


def highpass(x: np.ndarray, sr: int) -> np.ndarray:
    """Direct-form I biquad in fp32, output clamped to [-1, 1] after the recursion (torchaudio lfilter clamp=True)."""
    (b0, b1, b2), (a1, a2) = highpass_coeffs(sr)
    f = np.float32
    y = np.zeros(len(x), dtype=np.float32)
    x1 = x2 = y1 = y2 = f(0)
    for t in range(len(x)):
        xt = f(x[t])
        acc = f(f(f(b2) * x2) + f(f(b1) * x1)) + f(f(b0) * xt)
        acc = f(acc) - f(f(f(a2) * y2) + f(f(a1) * y1))
        yt = f(acc)
        y[t] = yt
        x2, x1, y2, y1 = x1, xt, y1, yt
    return np.clip(y, -1.0, 1.0)


def sinc_kernel_rows(n_rows: int = TARGET_LENGTH, orig_freq: int = 1, new_freq: int = TARGET_SR,
                     lowpass_filter_width: int = 6, rolloff: float = 0.99) -> torch.Tensor:
    """First `n_rows` polyphase filters of torchaudio's Hann-windowed sinc resampler for orig_freq -> new_freq, built
    with the same torch ops in the same order (bit-identical table), shape (n_rows, 2 * width + orig_freq).
    dataset.py:53-55 calls Resample(orig_freq=audio_segment.size(0), ...) and size(0) is the CHANNEL count (1), so the
    reference always resamples 1 Hz -> 16000 Hz and dataset.py:62-64 keeps the first 4000 outputs: those are filters
    0..3999 applied at input position 0."""
    base_freq = min(orig_freq, new_freq) * rolloff
    width = math.ceil(lowpass_filter_width * orig_freq / base_freq)
    idx = torch.arange(-width, width + orig_freq, dtype=torch.float64)[None, None] / orig_freq
    t = torch.arange(0, -new_freq, -1)[:, None, None] / new_freq + idx      # int64 / int -> fp32 quotient, as torchaudio
    t = t * base_freq
    t = t.clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base_freq / orig_freq
    kernels = torch.where(t == 0, torch.tensor(1.0).to(t), t.sin() / t)
    kernels = kernels * (window * scale)
    return kernels.to(torch.float32)[:n_rows, 0, :].contiguous(), width


def audio_window(waveform: np.ndarray, frame_end: int, total_frames: int, video_fps: float, sr: int) -> np.ndarray:
    """dataset.py:112-130 for a mono waveform (1-D fp32): returns input_values (4000,) fp32."""
    _, s, e = window_indices(frame_end, total_frames, video_fps, sr)
    seg = np.asarray(waveform[s:e], dtype=np.float32)
    if len(seg) == 0:
        raise ValueError("empty audio window (the reference divides by a NaN std here)")
    y = highpass(seg, sr)
    yt = torch.from_numpy(y)
    z = ((yt - yt.mean()) / yt.std()).numpy()                          # dataset.py:34-37 (unbiased std)
    K, width = sinc_kernel_rows()
    zp = np.zeros(K.shape[1], dtype=np.float32)                        # padded[k] = z[k - width]
    n = min(K.shape[1] - width, len(z))
    zp[width:width + n] = z[:n]
    o = (K.numpy().astype(np.float32) * zp[None, :]).sum(axis=1, dtype=np.float32)   # first 4000 resampled outputs
    m = o.mean(dtype=np.float32)
    v = o.var(dtype=np.float32)
    return ((o - m) / np.sqrt(v + np.float32(1e-7))).astype(np.float32)   # Wav2Vec2FeatureExtractor.zero_mean_unit_var_norm


def frame_to_tensor(frame_u8: np.ndarray) -> torch.Tensor:
    """train.py:70-75 without the Resize: ToTensor() (HWC uint8 -> CHW float / 255) then Normalize(0.5, 0.5)."""
    t = torch.from_numpy(frame_u8).permute(2, 0, 1).to(torch.float32).div(255)
    return (t - 0.5) / 0.5


# ----------------------------------------------------------------------------- literal call sites (pinning)
def reference_item(waveform_1xL: torch.Tensor, frame_end: int, total_frames: int, video_fps: float, sr: int):
    """dataset.py:98,112-130 line by line with the reference's own libraries (torchaudio, transformers)."""
    import torchaudio
    from transformers import Wav2Vec2FeatureExtractor
    output_frame_idx = min(frame_end, total_frames - 1)
    frame_duration = 1 / video_fps
    start_sec = max(0, (output_frame_idx - BUFFER_FRAMES) * frame_duration)
    end_sec = output_frame_idx * frame_duration
    start_sample = int(sr * start_sec)
    end_sample = int(sr * end_sec)
    audio_segment = waveform_1xL[:, start_sample:end_sample]
    audio_segment = torchaudio.functional.highpass_biquad(audio_segment, sr, 300)
    audio_segment = (audio_segment - audio_segment.mean()) / audio_segment.std()
    if audio_segment.size(0) != TARGET_SR:
        resampler = torchaudio.transforms.Resample(orig_freq=audio_segment.size(0), new_freq=TARGET_SR)
        # the full result has 16000 * L samples; only [:4000] survives dataset.py:62-64.  Resampling the first
        # 2 * width + 1 input samples gives the same first 4000 outputs (they depend on x[0..width] only) without
        # materialising 16000 * L floats.
        audio_segment = resampler(audio_segment[:, :16])
    if audio_segment.size(1) < TARGET_LENGTH:
        audio_segment = torch.cat([audio_segment, torch.zeros(audio_segment.size(0), TARGET_LENGTH - audio_segment.size(1))], dim=1)
    elif audio_segment.size(1) > TARGET_LENGTH:
        audio_segment = audio_segment[:, :TARGET_LENGTH]
    fe = Wav2Vec2FeatureExtractor(feature_size=1, sampling_rate=16000, padding_value=0.0, do_normalize=True,
                                  return_attention_mask=False)
    out = fe(audio_segment.squeeze(0).numpy(), return_tensors="pt", sampling_rate=TARGET_SR)
    return output_frame_idx, start_sample, end_sample, out["input_values"][0]
