"""Profiling driver (GPU): one forward + one backward attention call at N=32 (B2D_PROF_N overrides), L=4096, ch=64 (full-resolution layer shape)."""
import sys
import torch
sys.path.insert(0, ".")
from lipreading_video_generation_b200 import ops

g = torch.Generator(device="cuda").manual_seed(0)
import os
N, H, C = int(os.environ.get("B2D_PROF_N", "32")), 64, 64
qkv = torch.randn(N, H, H, 3 * C, device="cuda", generator=g).half()
d_o = torch.randn(N, H, H, C, device="cuda", generator=g).half()
for _ in range(2):
    out, lse = ops.attention_fwd(qkv, 1)
    dqkv = ops.attention_bwd(qkv, out, d_o, lse, 1)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), float(dqkv.float().abs().mean()))
