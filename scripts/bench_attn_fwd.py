"""Developer tool: time b2d_attention_fwd at the train64 / sample128 shapes (TFLOP/s, algorithmic 4 L^2 ch N)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
for N, S, ch in ((256, 64, 64), (32, 128, 64), (256, 32, 128)):
    qkv = (torch.randn(N, S, S, 3 * ch, device="cuda") * 1.5).half()
    for _ in range(2):
        out, lse = ops.attention_fwd(qkv, 1)
    ts = []
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.attention_fwd(qkv, 1)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    L = S * S
    # accuracy against an fp32 softmax on a few rows of image 0
    q, k, v = [t.float() for t in qkv[0].view(L, 3 * ch).split(ch, dim=1)]
    rows = torch.arange(0, L, max(1, L // 64), device="cuda")
    w = torch.softmax((q[rows] @ k.t()) / ch ** 0.5, dim=-1)
    ref = w @ v
    err = ((out[0].view(L, ch)[rows].float() - ref).norm() / ref.norm()).item()
    print(dict(N=N, L=L, ch=ch, ms=round(ms, 4), tflops=round(4.0 * L * L * ch * N / ms / 1e9, 1), rel_err=err), flush=True)
