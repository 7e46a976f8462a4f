"""Developer tool for ncu (round 2): one attention forward at the sample128 bench shape (N = 128, L = 16384, ch 64) and one
attention backward at the train128 bench shape (N = 256, L = 16384, ch 64), after a warm-up launch of each."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

S, ch = 128, 64
for N, do_bwd in ((128, False), (256, True)):
    qkv = (torch.randn(N, S, S, 3 * ch, device="cuda") * 1.5).half()
    for _ in range(2):
        out, lse = ops.attention_fwd(qkv, 1)
    if do_bwd:
        d_o = torch.randn(N, S, S, ch, device="cuda").half()
        for _ in range(2):
            ops.attention_bwd(qkv, out, d_o, lse, 1)
    torch.cuda.synchronize()
    del qkv
print("ok")
