"""Developer tool: time b2d_conv2d_wgrad on the weight-gradient shapes of the train64 step (N = 256 images), TFLOP/s
per shape and the class total.  B2D_WGRAD_HALO=0 selects the per-tap kernel for all shapes (A/B)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

# (H, C1, C2, Cout, count per step) of the 3x3 stride-1 convolutions of the train.py model at 64x64
SHAPES = [(64, 64, 0, 64, 7), (64, 256, 0, 64, 1), (64, 128, 64, 64, 1), (64, 64, 64, 64, 2), (64, 128, 0, 128, 1),
          (32, 64, 0, 128, 1), (32, 128, 0, 128, 6), (32, 256, 128, 128, 1), (32, 128, 128, 128, 1), (32, 128, 64, 128, 1),
          (32, 256, 0, 256, 1), (16, 128, 0, 256, 1), (16, 256, 0, 256, 10), (16, 256, 256, 256, 2), (16, 256, 128, 256, 1)]
N = int(os.environ.get("WG_N", "256"))
if os.environ.get("WG_ONLY"):                     # comma-separated shape indices (profiling runs)
    SHAPES = [SHAPES[int(i)] for i in os.environ["WG_ONLY"].split(",")]
out, total = [], 0.0
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
for H, C1, C2, Cout, cnt in SHAPES:
    x1 = torch.randn(N, H, H, C1, device="cuda").half()
    x2 = torch.randn(N, H, H, C2, device="cuda").half() if C2 else None
    dy = torch.randn(N, H, H, Cout, device="cuda").half()
    dw = torch.zeros(Cout, 9 * (C1 + C2), device="cuda")
    for _ in range(2):
        ops.conv2d_wgrad(x1, dy, dw, 3, 1, x2=x2)
    ts = []
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.conv2d_wgrad(x1, dy, dw, 3, 1, x2=x2)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    fl = 2.0 * 9 * (C1 + C2) * Cout * N * H * H
    out.append(dict(shape=[N, H, H, C1 + C2, Cout], count=cnt, ms=round(ms, 4), tflops=round(fl / ms / 1e9, 1)))
    total += ms * cnt
    print(out[-1], flush=True)
print(json.dumps(dict(halo=os.environ.get("B2D_WGRAD_HALO", "1"), class_ms_per_step=round(total, 3), shapes=out)))
