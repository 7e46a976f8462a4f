"""Profiling driver (GPU): GroupNorm(+SiLU) forward and backward at a full-resolution bench shape
(N=256, 64x64, C=128), plain and two-source (64+64 skip concat)."""
import sys
import torch
sys.path.insert(0, ".")
from lipreading_video_generation_b200 import ops

g = torch.Generator(device="cuda").manual_seed(0)
N, H, C = 256, 64, 128
x = torch.randn(N, H, H, C, device="cuda", generator=g).half()
dy = torch.randn(N, H, H, C, device="cuda", generator=g).half()
add = torch.randn(N, H, H, C, device="cuda", generator=g).half()
gamma = torch.rand(C, device="cuda", generator=g) + 0.5
beta = torch.randn(C, device="cuda", generator=g) * 0.1
dg, db = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
for _ in range(3):
    y, mean, rstd = ops.groupnorm_fwd(x, gamma, beta, act=True)
    dx1, _ = ops.groupnorm_bwd(x, gamma, beta, dy, mean, rstd, dg, db, act=True, add_a=add)
torch.cuda.synchronize()
e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
e0.record()
for _ in range(10):
    y, mean, rstd = ops.groupnorm_fwd(x, gamma, beta, act=True)
e1.record()
for _ in range(10):
    dx1, _ = ops.groupnorm_bwd(x, gamma, beta, dy, mean, rstd, dg, db, act=True, add_a=add)
e2.record()
torch.cuda.synchronize()
B = x.numel() * 2 / 1e9
print(f"fwd {e0.elapsed_time(e1) / 10:.4f} ms ({3 * B / (e0.elapsed_time(e1) / 10) * 1e3:.0f} GB/s of 3B)   "
      f"bwd {e1.elapsed_time(e2) / 10:.4f} ms ({6 * B / (e1.elapsed_time(e2) / 10) * 1e3:.0f} GB/s of 6B)")
