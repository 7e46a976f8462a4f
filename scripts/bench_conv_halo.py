"""3x3 stride-1 convolutions of the train64 step (N = 256): per-tap loads (B2D_CONV_HALO=0) against haloed patch tiles
(conv_gemm.cu, HALO variant), per epilogue variant.  CUDA events, three rotating input sets (> L2 at the 64x64 layers).
`python scripts/bench_conv_halo.py [reps]`; B2D_CONV_HALO_CFG=na,nb,ns overrides the ring depths of the BN = 128 path."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
N = 256
torch.manual_seed(0)
# (H, cin, cout, mode, launches per step)
SHAPES = [(64, 64, 64, "plain", 8), (64, 64, 64, "gnb", 8), (64, 64, 128, "gnb", 2), (64, 64, 192, "gnb", 1),
          (64, 64, 256, "plain", 1), (64, 128, 64, "stats", 2), (64, 256, 64, "stats", 1), (64, 192, 64, "stats", 1),
          (64, 128, 128, "stats", 2), (32, 64, 128, "plain", 1), (32, 128, 128, "stats", 7), (32, 128, 128, "gnb", 6),
          (32, 256, 256, "stats", 2), (32, 384, 128, "stats", 1), (32, 256, 128, "stats", 1), (32, 192, 128, "stats", 1),
          (32, 128, 384, "gnb", 1), (32, 128, 256, "gnb", 1), (32, 128, 192, "gnb", 1), (16, 256, 256, "stats", 10),
          (16, 256, 256, "gnb", 10), (16, 512, 256, "stats", 2), (16, 256, 512, "gnb", 2), (16, 384, 256, "stats", 1),
          (16, 256, 384, "gnb", 1)]


def setup(H, cin, cout, mode):
    sets = []
    w = ops.pack_conv_weight((torch.randn(cout, cin, 3, 3, device="cuda") / (cin * 9) ** 0.5))
    for _ in range(3):
        x = torch.randn(N, H, H, cin, device="cuda").half()
        res = torch.randn(N, H, H, cout, device="cuda").half() if mode != "gnb" else None
        bias = torch.randn(cout, device="cuda") if mode != "gnb" else None
        st = torch.zeros(N, cout // 2, 2, device="cuda", dtype=torch.float64) if mode == "stats" else None
        gnb = None
        if mode == "gnb":
            xg = torch.randn(N, H, H, cout, device="cuda").half()
            gnb = ops.GnBwdFuse(xg, None, torch.zeros(N, 32, device="cuda"), torch.ones(N, 32, device="cuda"),
                                torch.ones(cout, device="cuda"), torch.zeros(cout, device="cuda"), True,
                                torch.zeros(N, cout, 2, device="cuda"))
        out = torch.empty(N, H, H, cout, device="cuda", dtype=torch.float16)
        sets.append((x, res, bias, st, gnb, out))
    return w, sets


def timed(w, sets, cout):
    def f(i):
        x, res, bias, st, gnb, out = sets[i % 3]
        ops.conv2d(x, w, cout, 3, 1, bias=bias, residual=res, stats_out=st, gnb=gnb, out=out)
    for i in range(3):
        f(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        f(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


tot = [0.0, 0.0]
print(f"{'H':>3s} {'cin':>4s} {'cout':>4s} mode   x   per-tap ms   haloed ms   TFLOP/s (haloed)")
for H, cin, cout, mode, cnt in SHAPES:
    w, sets = setup(H, cin, cout, mode)
    t = []
    for h in ("0", "1"):
        os.environ["B2D_CONV_HALO"] = h
        t.append(timed(w, sets, cout))
    tot[0] += t[0] * cnt
    tot[1] += t[1] * cnt
    tf = 2.0 * N * H * H * cin * cout * 9 / (t[1] * 1e-3) / 1e12
    print(f"{H:3d} {cin:4d} {cout:4d} {mode:6s} {cnt:2d}   {t[0]:8.4f}   {t[1]:8.4f}   {tf:7.0f}")
    del w, sets
    torch.cuda.empty_cache()
print(f"class per step: per-tap {tot[0]:.3f} ms, haloed {tot[1]:.3f} ms")
