"""Profile target: one haloed-patch convolution (conv_gemm.cu, HALO variant).  `python scripts/prof_conv_halo.py H cin cout mode`;
under ncu: -k regex:conv_gemm_kernel -s 2 -c 1 --set full --import-source on."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

H, cin, cout = (int(v) for v in sys.argv[1:4])
mode = sys.argv[4] if len(sys.argv) > 4 else "plain"
N = 256
torch.manual_seed(0)
x = torch.randn(N, H, H, cin, device="cuda").half()
w = ops.pack_conv_weight(torch.randn(cout, cin, 3, 3, device="cuda") / (cin * 9) ** 0.5)
res = torch.randn(N, H, H, cout, device="cuda").half() if mode != "gnb" else None
bias = torch.randn(cout, device="cuda") if mode != "gnb" else None
st = torch.zeros(N, cout // 2, 2, device="cuda", dtype=torch.float64) if mode == "stats" else None
gnb = None
if mode == "gnb":
    xg = torch.randn(N, H, H, cout, device="cuda").half()
    gnb = ops.GnBwdFuse(xg, None, torch.zeros(N, 32, device="cuda"), torch.ones(N, 32, device="cuda"),
                        torch.ones(cout, device="cuda"), torch.zeros(cout, device="cuda"), True,
                        torch.zeros(N, cout, 2, device="cuda"))
for _ in range(4):
    ops.conv2d(x, w, cout, 3, 1, bias=bias, residual=res, stats_out=st, gnb=gnb)
torch.cuda.synchronize()
print("ok")
