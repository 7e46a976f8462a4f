"""Developer check (GPU): attention bwd and GroupNorm fwd/bwd vs fp32 torch autograd."""
import ctypes as C
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
sys.path.insert(0, "scripts")
from lipreading_video_generation_b200 import _lib as L  # noqa: E402
from dev_attn_check import ref_attn  # noqa: E402


def rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


def attn_bwd(N, Lq, heads, ch, new_order=0, time_it=False):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(5)
    Cc = heads * ch
    qkv = torch.randn(N, Lq, 3 * Cc, device=dev, generator=g).to(torch.float16)
    d_o = torch.randn(N, Lq, Cc, device=dev, generator=g).to(torch.float16)
    out = torch.empty(N, Lq, Cc, device=dev, dtype=torch.float16)
    lse = torch.empty(N, heads, Lq, device=dev)
    dqkv = torch.full((N, Lq, 3 * Cc), float("nan"), device=dev, dtype=torch.float16)
    lib = L.lib()
    delta = torch.empty(int(lib.b2d_attention_bwd_ws_floats(N, Lq, heads, ch)), device=dev)
    L.check(lib.b2d_attention_fwd(qkv.data_ptr(), 3 * Cc, N, Lq, heads, ch, new_order, out.data_ptr(), Cc,
                                  lse.data_ptr(), L.stream_ptr()), "fwd")

    def call():
        L.check(lib.b2d_attention_bwd(qkv.data_ptr(), 3 * Cc, out.data_ptr(), Cc, d_o.data_ptr(), Cc, lse.data_ptr(),
                                      N, Lq, heads, ch, new_order, dqkv.data_ptr(), 3 * Cc, delta.data_ptr(),
                                      L.stream_ptr()), "bwd")
    call()
    torch.cuda.synchronize()
    x = qkv.float().requires_grad_(True)
    ref = ref_attn(x, heads, ch, new_order)
    (gref,) = torch.autograd.grad(ref, x, d_o.float())
    e = rel(dqkv, gref)
    msg = f"attn bwd N={N} L={Lq} heads={heads} ch={ch} new={new_order}: rel err {e:.3e}"
    if time_it:
        for _ in range(3):
            call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        msg += f"  {ms:.3f} ms  {10.0 * Lq * Lq * ch * heads * N / ms / 1e9:.1f} TFLOP/s (5 GEMMs algorithmic)"
    print(msg, flush=True)
    return e


def gn(N, HW, C1, C2, act, with_add):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(7)
    Ct = C1 + C2
    x1 = (torch.randn(N, HW, C1, device=dev, generator=g) * 2 + 0.5).to(torch.float16)
    x2 = (torch.randn(N, HW, C2, device=dev, generator=g)).to(torch.float16) if C2 else None
    gamma = torch.randn(Ct, device=dev, generator=g); beta = torch.randn(Ct, device=dev, generator=g)
    dy = torch.randn(N, HW, Ct, device=dev, generator=g).to(torch.float16)
    add_a = torch.randn(N, HW, C1, device=dev, generator=g).to(torch.float16) if with_add else None
    a = L.GnArgs()
    a.x1, a.ld1, a.C1 = x1.data_ptr(), C1, C1
    a.x2, a.ld2, a.C2 = (x2.data_ptr() if C2 else None), C2, C2
    a.N, a.HW = N, HW
    a.gamma, a.beta, a.eps, a.act, a.drop_p, a.seed = gamma.data_ptr(), beta.data_ptr(), 1e-5, act, 0.0, 0
    lib = L.lib()
    wsb = lib.b2d_groupnorm_workspace_bytes(N, Ct)
    ws = torch.empty(wsb, device=dev, dtype=torch.uint8)
    y = torch.empty(N, HW, Ct, device=dev, dtype=torch.float16)
    mean = torch.empty(N, 32, device=dev); rstd = torch.empty(N, 32, device=dev)
    L.check(lib.b2d_groupnorm_fwd(C.byref(a), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), ws.data_ptr(), wsb,
                                  L.stream_ptr()), "gn fwd")
    dx1 = torch.empty_like(x1); dx2 = torch.empty_like(x2) if C2 else None
    dg = torch.zeros(Ct, device=dev); db = torch.zeros(Ct, device=dev)
    L.check(lib.b2d_groupnorm_bwd(C.byref(a), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx1.data_ptr(), C1,
                                  L.ptr(dx2), C2, L.ptr(add_a), C1, None, 0, 0, dg.data_ptr(), db.data_ptr(),
                                  ws.data_ptr(), wsb, L.stream_ptr()), "gn bwd")
    torch.cuda.synchronize()
    xin = (x1 if not C2 else torch.cat([x1, x2], -1)).float().requires_grad_(True)
    gp = gamma.clone().requires_grad_(True); bp = beta.clone().requires_grad_(True)
    ref = F.group_norm(xin.permute(0, 2, 1), 32, gp, bp, 1e-5)
    if act:
        ref = F.silu(ref)
    ref = ref.permute(0, 2, 1)
    gx, gg, gb = torch.autograd.grad(ref, (xin, gp, bp), dy.float())
    if with_add:
        gx = gx.clone(); gx[..., :C1] += add_a.float()
    errs = [rel(y, ref), rel(dx1, gx[..., :C1]), rel(dg, gg), rel(db, gb)]
    if C2:
        errs.append(rel(dx2, gx[..., C1:]))
    print(f"gn N={N} HW={HW} C=({C1},{C2}) act={act} add={with_add}: " + " ".join(f"{e:.2e}" for e in errs), flush=True)
    return max(errs)


if __name__ == "__main__":
    worst = 0.0
    worst = max(worst, gn(2, 256, 64, 0, 1, False))
    worst = max(worst, gn(3, 1024, 128, 64, 1, True))
    worst = max(worst, gn(2, 64, 256, 256, 0, True))
    worst = max(worst, gn(2, 4096, 192, 0, 1, False))
    worst = max(worst, attn_bwd(2, 256, 1, 64))
    worst = max(worst, attn_bwd(2, 1024, 1, 128))
    worst = max(worst, attn_bwd(2, 256, 1, 256))
    worst = max(worst, attn_bwd(2, 320, 2, 64))
    worst = max(worst, attn_bwd(1, 200, 2, 64, 1))
    worst = max(worst, attn_bwd(1, 64, 1, 256))
    attn_bwd(64, 4096, 1, 64, time_it=True)
    attn_bwd(256, 4096, 1, 64, time_it=True)     # bench shape (cfg2 full-res layer)
    attn_bwd(256, 1024, 1, 128, time_it=True)    # cfg2 half-res layer
    attn_bwd(64, 1024, 1, 128, time_it=True)
    attn_bwd(64, 256, 1, 256, time_it=True)
    print("WORST", worst)
    sys.exit(0 if worst < 2e-2 else 1)
