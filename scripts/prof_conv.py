"""Profiling driver (GPU): a few conv2d / wgrad / groupnorm launches at bench shapes (N=64 images of 64x64)."""
import sys
import torch
sys.path.insert(0, ".")
from lipreading_video_generation_b200 import ops

g = torch.Generator(device="cuda").manual_seed(0)
N, H = 64, 64
x = torch.randn(N, H, H, 64, device="cuda", generator=g).half()
res = torch.randn(N, H, H, 64, device="cuda", generator=g).half()
w1 = ops.pack_conv_weight(torch.randn(192, 64, 1, 1, device="cuda", generator=g) * 0.1)
w3 = ops.pack_conv_weight(torch.randn(64, 64, 3, 3, device="cuda", generator=g) * 0.05)
b192 = torch.randn(192, device="cuda", generator=g)
b64 = torch.randn(64, device="cuda", generator=g)
emb = torch.randn(N, 64, device="cuda", generator=g)
gamma, beta = torch.ones(64, device="cuda"), torch.zeros(64, device="cuda")
dy = torch.randn(N, H, H, 64, device="cuda", generator=g).half()
for _ in range(2):
    y1 = ops.conv2d(x, w1, 192, 1, 1, bias=b192)                       # qkv 1x1
    y3 = ops.conv2d(x, w3, 64, 3, 1, bias=b64, emb=emb, residual=res)  # ResBlock conv with fused epilogue
    dw = torch.zeros(64, 9 * 64, device="cuda")
    ops.conv2d_wgrad(x, dy, dw, 3, 1)
    yn, mean, rstd = ops.groupnorm_fwd(x, gamma, beta, act=True)
    dg, db = torch.zeros(64, device="cuda"), torch.zeros(64, device="cuda")
    ops.groupnorm_bwd(x, gamma, beta, dy, mean, rstd, dg, db, act=True, add_a=res)
torch.cuda.synchronize()
print("ok", float(y1.float().abs().mean()), float(y3.float().abs().mean()))
