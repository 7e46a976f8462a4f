"""Developer check (GPU): fused wav2vec2 layer 0 against the stock layer in float64, and kernel timings."""
import copy
import sys
import torch
sys.path.insert(0, ".")
from transformers import Wav2Vec2Config, Wav2Vec2Model  # noqa: E402
import lipreading_video_generation_b200.unet_audio as UA  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(0)
m = Wav2Vec2Model(Wav2Vec2Config())
with torch.no_grad():
    gn = m.feature_extractor.conv_layers[0].layer_norm
    gn.weight.uniform_(0.5, 1.5)
    gn.bias.uniform_(-0.5, 0.5)
stock32 = copy.deepcopy(m).cuda().feature_extractor.conv_layers[0]
stock64 = copy.deepcopy(m).double().cuda().feature_extractor.conv_layers[0]
fused = UA.patch_feature_layer0(copy.deepcopy(m)).cuda().feature_extractor.conv_layers[0]


def rel(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


for N, S in ((3, 4000), (70, 4000), (2, 1234)):
    a = torch.randn(N, 1, S, device="cuda")
    r = torch.randn(N, 512, (S - 10) // 5 + 1, device="cuda")
    outs = {}
    for name, layer, x, rr in (("stock32", stock32, a, r), ("stock64", stock64, a.double(), r.double()), ("fused", fused, a, r)):
        for p in layer.parameters():
            p.grad = None
        y = layer(x)
        (y * rr).sum().backward()
        outs[name] = [y] + [layer.get_parameter(k).grad for k in ("conv.weight", "layer_norm.weight", "layer_norm.bias")]
    for i, k in enumerate(("out", "d conv.weight", "d gn.weight", "d gn.bias")):
        print(f"N={N} S={S} {k:14s} fused-vs-f64 {rel(outs['fused'][i], outs['stock64'][i]):.2e}   "
              f"stock32-vs-f64 {rel(outs['stock32'][i], outs['stock64'][i]):.2e}")

a = torch.randn(256, 1, 4000, device="cuda")
for dt in (torch.float32, torch.bfloat16):
    for _ in range(2):
        for p in fused.parameters():
            p.grad = None
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        with torch.autocast("cuda", dtype=dt, enabled=dt != torch.float32):
            e[0].record()
            y = fused(a)
            e[1].record()
        g = torch.randn_like(y)
        e[1].record()
        y.backward(g)
        e[2].record()
        torch.cuda.synchronize()
    print(dt, f"fwd {e[0].elapsed_time(e[1]):.3f} ms  bwd {e[1].elapsed_time(e[2]):.3f} ms")
