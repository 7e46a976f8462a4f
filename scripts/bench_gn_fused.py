"""Developer tool: time the one-pass GroupNorm kernels the engine uses (b2d_groupnorm_apply from pair statistics;
b2d_groupnorm_bwd_apply from the fused data-gradient epilogue's sums) on the train64 shapes; GB/s algorithmic."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

N = 256
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)


def timeit(fn):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(6):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


tot_f = tot_b = 0.0
for HW, Cc, cnt in ((4096, 64, 13), (1024, 128, 11), (256, 256, 16), (4096, 128, 2), (4096, 192, 1), (1024, 384, 1)):
    H = int(HW ** 0.5)
    x = torch.randn(N, H, H, Cc, device="cuda").half()
    dyh = torch.randn(N, H, H, Cc, device="cuda").half()
    add = torch.randn(N, H, H, Cc, device="cuda").half()
    gamma, beta = torch.ones(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    st = ops.pair_stats(x)
    y, mean, rstd, _ = ops.groupnorm_apply(x, st, gamma, beta, act=True)
    ab = torch.randn(N, Cc, 2, device="cuda")
    dg, db = torch.zeros(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    cs = torch.zeros(N, Cc, device="cuda")
    ovf = torch.zeros(4, device="cuda", dtype=torch.int32)
    f = timeit(lambda: ops.groupnorm_apply(x, st, gamma, beta, act=True))
    fd = timeit(lambda: ops.groupnorm_apply(x, st, gamma, beta, act=True, drop_p=0.1, seed=3, want_mask=True))
    b = timeit(lambda: ops.groupnorm_bwd_apply(x, gamma, beta, dyh, mean, rstd, ab, dg, db, add_a=add, colsum_out=cs, overflow=ovf))
    nb = x.numel() * 2
    print(dict(shape=[N, HW, Cc], fwd_ms=round(f, 4), fwd_gbs=round(2 * nb / f / 1e6), fwd_drop_ms=round(fd, 4),
               bwd_ms=round(b, 4), bwd_gbs=round(4 * nb / b / 1e6)), flush=True)
    tot_f += f * cnt
    tot_b += b * cnt
print(json.dumps(dict(bps_f=os.environ.get("B2D_GN_BPS_F"), bps_b=os.environ.get("B2D_GN_BPS_B"), fwd_class_ms=round(tot_f, 3),
                      bwd_class_ms=round(tot_b, 3))))
