"""TEST INFRASTRUCTURE ONLY: generates tests/golden/audio_window.pt from the reference's call sites evaluated with the
reference's own libraries (oracle/audio_window.reference_item: torchaudio highpass_biquad / Resample, HF feature
extractor; torchvision ToTensor / Normalize) and asserts the plain restatement agrees.  Run in the build container:
    python -m oracle.make_golden_audio
"""
import os

import numpy as np
import torch

from oracle import audio_window as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# (video, frame_end) cases: start clamped to 0 (frame_end < 5), frame_end beyond the clip (min with total - 1),
# non-integer fps, 48 kHz track, window running past the end of a short track (Python slice clipping)
VIDEOS = [dict(name="a.mp4", T=40, fps=25.0, sr=16000, L=27000),
          dict(name="b.mp4", T=75, fps=29.97, sr=16000, L=41000),
          dict(name="c.mp4", T=30, fps=23.976, sr=48000, L=50000),
          dict(name="d.mp4", T=50, fps=30.0, sr=16000, L=20000)]
ITEMS = [("a.mp4", 30), ("a.mp4", 3), ("a.mp4", 999), ("b.mp4", 50), ("b.mp4", 74), ("c.mp4", 17), ("c.mp4", 7),
         ("d.mp4", 40), ("d.mp4", 6), ("b.mp4", 5)]
H = W = 16


def main():
    import torchvision.transforms as T
    g = torch.Generator().manual_seed(20261105)
    vids = {}
    for v in VIDEOS:
        wave = torch.randn(1, v["L"], generator=g) * 0.2 + 0.05 * torch.sin(torch.arange(v["L"]) * 0.01)[None]
        frames = torch.randint(0, 256, (v["T"], H, W, 3), generator=g, dtype=torch.uint8)
        vids[v["name"]] = dict(v, wave=wave, frames=frames)
    tf = T.Compose([T.ToTensor(), T.Normalize(mean=[0.5, 0.5, 0.5], std=[0.5, 0.5, 0.5])])     # train.py:70-75 minus Resize
    idx, audio, fin, fout = [], [], [], []
    for name, fe in ITEMS:
        v = vids[name]
        i, s, e, iv = A.reference_item(v["wave"], fe, v["T"], v["fps"], v["sr"])
        idx.append((i, s, e)); audio.append(iv)
        fin.append(tf(v["frames"][0].numpy())); fout.append(tf(v["frames"][i].numpy()))
        # the restatement agrees with the reference call sites
        assert (i, s, e) == A.window_indices(fe, v["T"], v["fps"], v["sr"])
        o = A.audio_window(v["wave"][0].numpy(), fe, v["T"], v["fps"], v["sr"])
        assert np.abs(o - iv.numpy()).max() <= 1e-4, np.abs(o - iv.numpy()).max()   # fp32 chain; north_star fp32 bar 1e-4
        assert torch.equal(A.frame_to_tensor(v["frames"][i].numpy()), fout[-1])
    out = dict(videos=[dict(name=v["name"], T=v["T"], fps=v["fps"], sr=v["sr"], wave=vids[v["name"]]["wave"],
                            frames=vids[v["name"]]["frames"]) for v in VIDEOS],
               items=ITEMS, indices=torch.tensor(idx, dtype=torch.int64), input_values=torch.stack(audio),
               input_frame=torch.stack(fin), output_frame=torch.stack(fout))
    path = os.path.join(ROOT, "tests", "golden", "audio_window.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes; indices", idx)


if __name__ == "__main__":
    main()
