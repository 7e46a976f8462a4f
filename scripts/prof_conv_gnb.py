"""Profile target: the default (per-tap) convolution kernel with the GroupNorm-backward epilogue.
`python scripts/prof_conv_gnb.py H cin cout k`; under ncu: -k regex:conv_gemm_kernel -s 2 -c 1 --set full --import-source on."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipreading_video_generation_b200 import ops

H, cin, cout, k = (int(v) for v in sys.argv[1:5])
N = 256
torch.manual_seed(0)
x = torch.randn(N, H, H, cin, device="cuda").half()
w = ops.pack_conv_weight(torch.randn(cout, cin, k, k, device="cuda") / (cin * k * k) ** 0.5)
xg = torch.randn(N, H, H, cout, device="cuda").half()
gnb = ops.GnBwdFuse(xg, None, torch.zeros(N, 32, device="cuda"), torch.ones(N, 32, device="cuda"),
                    torch.ones(cout, device="cuda"), torch.zeros(cout, device="cuda"), True,
                    torch.zeros(N, cout, 2, device="cuda"))
for _ in range(4):
    ops.conv2d(x, w, cout, k, 1, gnb=gnb)
torch.cuda.synchronize()
print("ok")
