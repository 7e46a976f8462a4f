"""Developer check (GPU): wav2vec2-base encode (fwd + bwd, bf16 autocast) with the stock and the short-sequence
positional conv / waveform conv, at the train64 bench shape (256 windows of 4000 samples)."""
import os
import sys
import torch
sys.path.insert(0, ".")
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
from lipreading_video_generation_b200.unet_audio import Wav2Vec2Encoder  # noqa: E402

torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
enc = Wav2Vec2Encoder().cuda().train()
pce = enc.wav2vec2.encoder.pos_conv_embed
c0 = enc.wav2vec2.feature_extractor.conv_layers[0].conv
fast, stock = (pce.forward, c0.forward), (pce._b2d_stock_forward, c0._b2d_stock_forward)
a = torch.randn(256, 4000, device="cuda")


def run():
    with torch.autocast("cuda", dtype=torch.bfloat16):
        f = enc({"input_values": a}).mean(dim=1)
    f.float().square().sum().backward()


for name, fwd in (("stock", stock), ("short-sequence", fast), ("stock", stock), ("short-sequence", fast)):
    pce.forward, c0.forward = fwd
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:16s} {e0.elapsed_time(e1) / 5:.3f} ms per fwd+bwd", flush=True)
