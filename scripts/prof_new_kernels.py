"""Profiling driver (GPU): the kernels added for rows N2 / N3 and the one-launch operand refresh at bench-like sizes.
ncu --set full -k regex:'adam_multi|audio_window|frame_gather|pack_conv_weights_multi' python scripts/prof_new_kernels.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lipreading_video_generation_b200 as P

dev = "cuda"
# ---- Adam: 110.7 M parameters in 344 tensors (the train.py model's count), one launch
g = torch.Generator().manual_seed(0)
sizes = [94_371_840 // 300] * 300 + [16_350_000 // 44] * 44
params = [torch.nn.Parameter(torch.randn(n, device=dev)) for n in sizes]
opt = P.Adam(params, lr=1e-2)
for _ in range(2):
    for p in params:
        p.grad = torch.randn_like(p)
    opt.step()
# ---- input pipeline: 256 items (one train64 batch) from 16 clips of 75 frames 64x64 and 3 s of 16 kHz audio
st = P.DeviceVideoStore(dev)
for v in range(16):
    st.add(f"v{v}", torch.randint(0, 256, (75, 64, 64, 3), generator=g, dtype=torch.uint8), torch.randn(1, 48000, generator=g) * 0.1, 25.0, 16000)
items = [P.FrameItem(f"v{i % 16}", 0, 5 + (i * 7) % 70) for i in range(256)]
b = P.TalkingFaceFrameBatcher(st, items)
for _ in range(2):
    fin, fout, audio = b.batch(list(range(256)))
b.check()
# ---- operand refresh of the train.py torso (166 descriptors)
m = P.UNetModel(image_size=64, in_channels=195, model_channels=64, out_channels=3, num_res_blocks=2,
                attention_resolutions=(1, 2, 4), dropout=0.0, channel_mult=(1, 2, 4)).to(dev).train()
x = torch.randn(2, 195, 64, 64, device=dev); t = torch.tensor([1, 500], device=dev)
for _ in range(3):
    out = m(x, t)
    out.square().mean().backward()
torch.cuda.synchronize()
print("ok", float(audio["input_values"].abs().mean()), float(out.abs().mean()))
