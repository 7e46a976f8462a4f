"""Developer tool: time the two-kernel GroupNorm backward (b2d_groupnorm_bwd: reduce + apply + parameter sums) on the
train64 shapes with the options the engine uses toggled one by one -- which of SiLU' (MUFU), the residual add, the column
sums or the dropout mask costs what; GB/s are algorithmic bytes (x, dy read twice, adds read once, dx written)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

N = 256
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
rows = []
for HW, Cc in ((4096, 64), (1024, 128), (4096, 128), (256, 256)):
    H = int(HW ** 0.5)
    x = torch.randn(N, H, H, Cc, device="cuda").half()
    dy = torch.randn(N, H, H, Cc, device="cuda").half()
    add = torch.randn(N, H, H, Cc, device="cuda").half()
    gamma, beta = torch.ones(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    _, mean, rstd = ops.groupnorm_fwd(x, gamma, beta)
    dg, db = torch.zeros(Cc, device="cuda"), torch.zeros(Cc, device="cuda")
    cs = torch.zeros(N, Cc, device="cuda")
    ovf = torch.zeros(4, device="cuda", dtype=torch.int32)
    for name, kw in (("plain act=0", dict(act=False)), ("act", dict(act=True)), ("act+add_a", dict(act=True, add_a=add)),
                     ("act+add_a+colsum", dict(act=True, add_a=add, colsum_out=cs)),
                     ("act+add_a+colsum+ovf", dict(act=True, add_a=add, colsum_out=cs, overflow=ovf)),
                     ("act+drop+add_a+colsum+ovf", dict(act=True, add_a=add, colsum_out=cs, overflow=ovf, drop_p=0.1, seed=5))):
        ts = []
        for _ in range(6):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.groupnorm_bwd(x, gamma, beta, dy, mean, rstd, dg, db, **kw)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        nbytes = x.numel() * 2 * (5 + (1 if "add_a" in kw else 0))
        rows.append(dict(shape=[N, HW, Cc], variant=name, ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 0)))
        print(rows[-1], flush=True)
print(json.dumps(rows))
