"""Profile target: conv_gemm_kernel epilogue variants (plain / statistics / GroupNorm-backward) on the epilogue-bound
shapes of the train64 step.  `python scripts/prof_conv_epi.py [reps]` prints CUDA-event times; run under ncu with
-k regex:conv_gemm for the source-level view."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
g = torch.Generator().manual_seed(0)
N = 256


def run(H, cin, cout, k, mode):
    x = torch.randn(N, H, H, cin, generator=g).half().cuda()
    w = ops.pack_conv_weight((torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).cuda())
    res = torch.randn(N, H, H, cout, generator=g).half().cuda() if mode != "gnb" else None
    bias = torch.randn(cout).cuda() if mode != "gnb" else None
    st = torch.zeros(N, cout // 2, 2, device="cuda", dtype=torch.float64) if mode == "stats" else None
    gnb = None
    if mode == "gnb":
        xg = torch.randn(N, H, H, cout, generator=g).half().cuda()
        gnb = ops.GnBwdFuse(xg, None, torch.zeros(N, 32).cuda(), torch.ones(N, 32).cuda(), torch.ones(cout).cuda(),
                            torch.zeros(cout).cuda(), True, torch.zeros(N, cout, 2).cuda())
    f = lambda: ops.conv2d(x, w, cout, k, 1, bias=bias, residual=res, stats_out=st, gnb=gnb)   # noqa: E731
    for _ in range(2):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        f()
    e1.record()
    torch.cuda.synchronize()
    print(f"H={H} cin={cin} cout={cout} k={k} {mode:6s} {e0.elapsed_time(e1) / reps:.4f} ms")


for shape in ((64, 64, 64, 1), (64, 64, 64, 3), (32, 128, 128, 3)):
    for mode in ("plain", "stats", "gnb"):
        run(*shape, mode)
