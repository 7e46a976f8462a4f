"""Device-side batches of `TalkingFaceFrameDataset` items (SURVEY 8f row N2;
/root/reference/video-generation/dataset.py:78-136, transforms of diffusion/train.py:70-75).

The reference builds every item on the host: decode two frames, load the whole audio track, slice the 5-frame audio
window, high-pass, normalise, "resample", processor-normalise, and the DataLoader collates.  Here decoded clips
(uint8 frames + mono waveform) are kept resident in HBM by `DeviceVideoStore` and `TalkingFaceFrameBatcher.batch()`
produces the collated batch `(input_frame, output_frame, {"input_values": (B, 1, 4000)})` with two kernel launches
(`b2d_audio_window`, `b2d_frame_gather`): the frame / audio-sample index arithmetic is done on the device in IEEE
double, bit-exact with the reference's Python floats; the float chain follows the reference's operation order.
Video decoding (decord), audio demuxing (torchaudio.load) and the PIL Resize stay host I/O and are out of scope:
frames are stored at the model resolution.  No CPU fallback."""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib as L

BUFFER_FRAMES = 5           # dataset.py:28
TARGET_LENGTH = 4000        # dataset.py:125
TARGET_SR = 16000           # dataset.py:126


class FrameItem:
    """dataset.py:44-48."""

    def __init__(self, video_path, frame_start, frame_end):
        self.video_path = video_path
        self.frame_start = frame_start
        self.frame_end = frame_end


def sinc_rows_1hz_to_16k(n_rows: int = TARGET_LENGTH) -> torch.Tensor:
    """dataset.py:53-55 builds `Resample(orig_freq=audio_segment.size(0), new_freq=16000)` and size(0) is the channel
    count, 1; dataset.py:62-64 keeps the first 4000 of the 16000 * L outputs.  Those are polyphase filters 0..3999 of
    torchaudio's Hann-windowed sinc kernel (lowpass_filter_width 6, rolloff 0.99 -> 15 taps) applied at input
    position 0.  The table is built with the same torch operations in the same order as torchaudio does."""
    orig_freq, new_freq, lowpass_filter_width, rolloff = 1, TARGET_SR, 6, 0.99
    base_freq = min(orig_freq, new_freq) * rolloff
    width = math.ceil(lowpass_filter_width * orig_freq / base_freq)
    idx = torch.arange(-width, width + orig_freq, dtype=torch.float64)[None, None] / orig_freq
    t = torch.arange(0, -new_freq, -1)[:, None, None] / new_freq + idx
    t = t * base_freq
    t = t.clamp_(-lowpass_filter_width, lowpass_filter_width)
    window = torch.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base_freq / orig_freq
    kernels = torch.where(t == 0, torch.tensor(1.0).to(t), t.sin() / t)
    kernels = kernels * (window * scale)
    return kernels.to(torch.float32)[:n_rows, 0, :].contiguous()


def highpass_biquad_coeffs(sr: int, cutoff_freq: float = 300.0, Q: float = 0.707):
    """(b0, b1, b2, a1, a2) / a0 in fp32 as torchaudio.functional.highpass_biquad forms them (dataset.py:30-32)."""
    w0 = 2 * math.pi * cutoff_freq / sr
    alpha = math.sin(w0) / 2.0 / Q
    t = torch.tensor([(1 + math.cos(w0)) / 2, -1 - math.cos(w0), (1 + math.cos(w0)) / 2, 1 + alpha, -2 * math.cos(w0),
                      1 - alpha], dtype=torch.float32)
    b = t[:3] / t[3]
    a = t[3:] / t[3]
    return [float(b[0]), float(b[1]), float(b[2]), float(a[1]), float(a[2])]


class DeviceVideoStore:
    """Decoded clips resident in HBM: all frames in one uint8 (sum T, H, W, 3) tensor, all mono tracks in one fp32
    tensor, with per-video offsets."""

    def __init__(self, device="cuda"):
        self.device = torch.device(device)
        self._frames, self._waves, self.meta, self.index = [], [], [], {}
        self.frames = self.waves = None

    def add(self, video_path, frames_u8: torch.Tensor, waveform: torch.Tensor, video_fps: float, sample_rate: int):
        if frames_u8.dtype != torch.uint8 or frames_u8.dim() != 4 or frames_u8.shape[3] != 3:
            raise ValueError("frames must be uint8 (T, H, W, 3) as decord's asnumpy() returns them")
        if video_fps == 0:
            raise ValueError("FPS is zero, which may indicate an issue with the video file or codec.")   # dataset.py:91-92
        wave = waveform.reshape(-1) if waveform.dim() == 1 or waveform.shape[0] == 1 else None
        if wave is None:
            raise ValueError("only mono tracks (1, L) are supported")
        self.index[video_path] = len(self.meta)
        self.meta.append((int(frames_u8.shape[0]), float(video_fps), int(sample_rate), int(wave.numel())))
        self._frames.append(frames_u8)
        self._waves.append(wave.to(torch.float32))
        self.frames = None

    def finalize(self):
        if self.frames is None:
            hw = {tuple(f.shape[1:3]) for f in self._frames}
            if len(hw) != 1:
                raise ValueError("all stored clips must share one resolution (the Resize of train.py:72 is host-side)")
            self.frames = torch.cat(self._frames).contiguous().to(self.device)
            self.waves = torch.cat(self._waves).contiguous().to(self.device)
            t = np.array([m[0] for m in self.meta], dtype=np.int64)
            n = np.array([m[3] for m in self.meta], dtype=np.int64)
            self.first_frame = torch.from_numpy(np.cumsum(t) - t).to(self.device)
            self.wave_offset = torch.from_numpy(np.cumsum(n) - n).to(self.device)
            self.wave_len = torch.from_numpy(n).to(self.device)
        return self


class TalkingFaceFrameBatcher:
    """`TalkingFaceFrameDataset(frame_items, ...)` + DataLoader collation for a list of item indices."""

    def __init__(self, store: DeviceVideoStore, frame_items, frame_rate=30):
        self.store = store.finalize()
        self.frame_items = list(frame_items)
        self.frame_rate = frame_rate
        self.sinc_rows = sinc_rows_1hz_to_16k().to(store.device)
        self._biquad = {}
        self.err_flag = torch.zeros(1, dtype=torch.int32, device=store.device)
        self.last_indices = None

    def __len__(self):
        return len(self.frame_items)

    def host_tables(self, indices):
        """Per-item launch arguments (host logic, testable without a GPU)."""
        vid, fend, tot, fps, sr, bq = [], [], [], [], [], []
        for i in indices:
            it = self.frame_items[i]
            v = self.store.index[it.video_path]
            t, f, s, _ = self.store.meta[v]
            vid.append(v); fend.append(int(it.frame_end)); tot.append(t); fps.append(f); sr.append(s)
            if s not in self._biquad:
                self._biquad[s] = highpass_biquad_coeffs(s)
            bq.append(self._biquad[s])
        return (np.array(vid, np.int32), np.array(fend, np.int64), np.array(tot, np.int64), np.array(fps, np.float64),
                np.array(sr, np.int32), np.array(bq, np.float32).reshape(-1, 5))

    def batch(self, indices):
        """-> (input_frame (B,3,H,W), output_frame (B,3,H,W), {"input_values": (B,1,4000)}) on the device."""
        st = self.store
        if not st.frames.is_cuda:
            raise RuntimeError("TalkingFaceFrameBatcher runs on sm_100a kernels only (no CPU fallback)")
        B = len(indices)
        if B == 0:
            H, W = st.frames.shape[1:3]
            z = torch.empty(0, 3, H, W, device=st.device, dtype=torch.float32)
            self.last_indices = torch.empty(0, 3, device=st.device, dtype=torch.int64)
            return z, z.clone(), {"input_values": torch.empty(0, 1, TARGET_LENGTH, device=st.device, dtype=torch.float32)}
        dev = [torch.from_numpy(a).pin_memory().to(st.device, non_blocking=True) for a in self.host_tables(indices)]
        vid, fend, tot, fps, sr, bq = dev
        audio = torch.empty(B, 1, TARGET_LENGTH, device=st.device, dtype=torch.float32)
        idx = torch.empty(B, 3, device=st.device, dtype=torch.int64)
        lib = L.lib()
        L.check(lib.b2d_audio_window(st.waves.data_ptr(), st.wave_offset.data_ptr(), st.wave_len.data_ptr(), vid.data_ptr(),
                                     fend.data_ptr(), tot.data_ptr(), fps.data_ptr(), sr.data_ptr(),
                                     self.sinc_rows.data_ptr(), bq.data_ptr(), B, audio.data_ptr(), idx.data_ptr(),
                                     self.err_flag.data_ptr(), L.stream_ptr()), "b2d_audio_window")
        H, W = st.frames.shape[1:3]
        fr = torch.empty(B, 2, 3, H, W, device=st.device, dtype=torch.float32)
        L.check(lib.b2d_frame_gather(st.frames.data_ptr(), st.first_frame.data_ptr(), vid.data_ptr(), idx.data_ptr(), B, H, W,
                                     fr.data_ptr(), L.stream_ptr()), "b2d_frame_gather")
        self.last_indices = idx          # (output_frame_idx, start_sample, end_sample) per item
        return fr[:, 0], fr[:, 1], {"input_values": audio}

    def check(self):
        """One device->host read: raises if any window so far was empty (the reference yields NaN audio and its
        `except` returns (None, None)) or longer than the kernel's shared-memory window."""
        e = int(self.err_flag.item())
        if e:
            self.err_flag.zero_()
            raise ValueError("empty audio window" if e == 1 else
                             f"audio window longer than {L.lib().b2d_audio_window_max_samples()} samples")
