import sys, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import os
os.environ["B2D_WAV2VEC2_RANDOM_INIT"] = "1"
import lipreading_video_generation_b200 as P
from test_gpu_unet import build_tiny, rel
g = torch.load("tests/golden/unet_tiny.pt", weights_only=False)
model = build_tiny(g)
sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
x0, eps, t = g["x0"].cuda(), g["eps"].cuda(), g["t"].cuda()
xt = sched.add_noise(x0, eps, t)
pred = model(xt, g["cond"].cuda(), {"feats": g["feats"].cuda()}, t)
print("pred rel", rel(pred, g["pred"]))
loss = P.MSELoss()(pred, eps); print("loss", loss.item(), g["loss"].item())
loss.backward()
named = dict(model.named_parameters())
for k, r in g["grads"].items():
    print(f"{k:50s} {rel(named[k].grad, r):.4f}")
errs = sorted(((abs(named[k].grad.float().norm().item() - n) / (n + 1e-12)), k) for k, n in g["grad_norms"].items())
print("worst norm errs", errs[-5:])
