"""Developer tool for ncu: the HBM-bound kernels north_star names at the BASELINE configs[2] / [3] sizes -- q_sample, MSE
forward / backward (256 images 128x128), the DDPM / DDIM sampler updates (128 images), multi-tensor Adam (110.7 M
parameters), one-pass GroupNorm forward / backward (256 x 4096 x 64) -- three launches each."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
import torch
import lipreading_video_generation_b200 as P
from lipreading_video_generation_b200 import ops

dev = torch.device("cuda")
sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
crit = P.MSELoss()
x0 = torch.rand(256, 3, 128, 128, device=dev); eps = torch.randn_like(x0)
t = torch.randint(0, 1000, (256,), device=dev)
pred = torch.randn_like(x0).requires_grad_(True)
xt = torch.randn(128, 3, 128, 128, device=dev); e = torch.randn_like(xt); z = torch.randn_like(xt)
params = [torch.nn.Parameter(torch.randn(n, device=dev)) for n in (94_400_000, 16_300_000)]
for p in params:
    p.grad = torch.randn_like(p)
opt = P.Adam(params, lr=1e-4)
x = torch.randn(256, 64, 64, 64, device=dev).half()
dyh = torch.randn_like(x); add = torch.randn_like(x)
gamma, beta = torch.ones(64, device=dev), torch.zeros(64, device=dev)
st = ops.pair_stats(x)
y, mean, rstd, _ = ops.groupnorm_apply(x, st, gamma, beta, act=True)
ab = torch.randn(256, 64, 2, device=dev)
dg, db = torch.zeros(64, device=dev), torch.zeros(64, device=dev)
for _ in range(3):
    with torch.no_grad():
        sched.add_noise(x0, eps, t)
        sched.sample_prev_timestep(xt, e, 500, z=z)
        sched.ddim_step(xt, e, 500, 480)
    loss = crit(pred, eps)
    loss.backward()
    opt.step()
    ops.groupnorm_apply(x, st, gamma, beta, act=True)
    ops.groupnorm_bwd_apply(x, gamma, beta, dyh, mean, rstd, ab, dg, db, add_a=add)
torch.cuda.synchronize()
print("ok")
