"""Developer trace (GPU, build with B2D_EXTRA_NVCC_FLAGS=-DFWD2_TRACE): clock64 stamps of the phases of attn_fwd2_kernel's
CTA (0,0,0): softmax warp of group 0 / group 1 and the tcgen05 issuer, per KV step."""
import os
import sys
sys.path.insert(0, ".")
import torch
buf = torch.zeros(17 * 64 * 8, dtype=torch.int64, device="cuda")
os.environ["B2D_FWD2_TRACE_PTR"] = str(buf.data_ptr())
from lipreading_video_generation_b200 import ops  # noqa: E402
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
g = torch.Generator(device="cuda").manual_seed(0)
qkv = torch.randn(N, 64, 64, 192, device="cuda", generator=g).half()
for _ in range(3):
    out, lse = ops.attention_fwd(qkv, 1)
torch.cuda.synchronize()
t = buf.cpu().view(17, 64, 8)
base = int(t[16, 0, 0])
names = ["top", "s_full", "ld+s_free", "max+xchg", "pv_done", "rescale?", "exp", "P+p_full"]
for j in (12, 13, 14):
    print(f"---- step {j}: per softmax warp (group, lane quarter -> SMSP = (2 + 4*group + qd) % 4), stamps relative to the issuer's step-0 top")
    print("            " + " ".join(f"{n:>9s}" for n in names) + "   | phase durations")
    for role in range(16):
        r = [int(v) - base for v in t[role, j]]
        print(f"  g{role // 8} half{(role % 8) // 4} smsp{(2 + role) % 4} " + " ".join(f"{v:9d}" for v in r) + "   | " +
              " ".join(f"{r[k + 1] - r[k]:5d}" for k in range(7)) + f"  period {r[0] - int(t[role, j - 1, 0]) + base:5d}")
    r = [int(v) - base for v in t[16, j]]
    print("  issuer [top0 s_free0 pre_p0 p_full0 | top1 s_free1 pre_p1 p_full1] " + " ".join(f"{v:9d}" for v in r))
