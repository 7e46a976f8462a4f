"""Developer experiment (GPU): do the feature-extractor conv layers 1..6 of HF wav2vec2 run without cuDNN's
nchw<->nhwc conversion kernels when layer 0 hands them a channels-last (B, C, T) tensor (memory (B, T, C))?"""
import collections
import re
import sys
import torch
sys.path.insert(0, ".")
from transformers import Wav2Vec2Config, Wav2Vec2Model  # noqa: E402
from torch.profiler import profile, ProfilerActivity  # noqa: E402

torch.backends.cudnn.benchmark = True
m = Wav2Vec2Model(Wav2Vec2Config()).cuda().train()
layers = m.feature_extractor.conv_layers[1:]


def run(x):
    with torch.autocast("cuda", dtype=torch.bfloat16):
        h = x
        for l in layers:
            h = l(h)
        h = h.transpose(1, 2)
        h = m.feature_projection(h)[0]
    h.float().square().sum().backward()
    return h


for name, mk in (("contiguous (B,C,T)", lambda: torch.randn(256, 512, 799, device="cuda", dtype=torch.bfloat16)),
                 ("channels-last strides", lambda: torch.randn(256, 799, 512, device="cuda", dtype=torch.bfloat16).transpose(1, 2))):
    x = mk().requires_grad_(True)
    for _ in range(3):
        run(x)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        h = run(x)
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0, 0.0])
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA:
            k = re.sub(r"\(.*", "", re.sub(r"<.*", "", e.name))
            agg[k][0] += 1
            agg[k][1] += e.time_range.end - e.time_range.start
    tot = sum(v for _, v in agg.values())
    print(f"== {name}: busy {tot / 1e3:.3f} ms, x.grad strides {x.grad.stride()}, out strides {h.stride()}")
    for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:9]:
        print(f"   {v / 1e3:7.3f} ms x{n:3d} {k[:100]}")
