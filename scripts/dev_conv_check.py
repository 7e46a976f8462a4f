"""Developer check (GPU): conv2d / wgrad kernels against torch fp32 on fp16-rounded inputs."""
import ctypes as C
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from lipreading_video_generation_b200 import _lib as L  # noqa: E402


def pack_w(w):  # OIHW fp32 -> [CoutPad][tap][Cin] bf16
    co, ci, kh, kw = w.shape
    cop = (co + 15) // 16 * 16
    p = torch.zeros(cop, kh * kw * ci, device=w.device, dtype=torch.float16)
    p[:co] = w.permute(0, 2, 3, 1).reshape(co, -1).to(torch.float16)
    return p, cop


def run_conv(N, H, W, C1, C2, Cout, k, stride, extras=True):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1)
    x = torch.randn(N, H, W, C1, device=dev, generator=g).to(torch.float16)
    x2 = torch.randn(N, H, W, C2, device=dev, generator=g).to(torch.float16) if C2 else None
    w = torch.randn(Cout, C1 + C2, k, k, device=dev, generator=g) * 0.05
    wp, cop = pack_w(w)
    Ho, Wo = H // stride, W // stride
    bias = torch.randn(Cout, device=dev, generator=g) if extras else None
    emb = torch.randn(N, Cout, device=dev, generator=g) if extras else None
    res = torch.randn(N, Ho, Wo, Cout, device=dev, generator=g).to(torch.float16) if extras else None
    out = torch.empty(N, Ho, Wo, Cout, device=dev, dtype=torch.float16)
    a = L.ConvArgs()
    a.x, a.ldx, a.C1 = x.data_ptr(), C1, C1
    a.x2, a.ldx2, a.C2 = (x2.data_ptr() if C2 else None), C2, C2
    a.N, a.H, a.W, a.ksize, a.stride = N, H, W, k, stride
    a.w, a.Cout, a.CoutPad = wp.data_ptr(), Cout, cop
    a.bias = L.ptr(bias); a.emb = L.ptr(emb); a.ldemb = Cout; a.residual = L.ptr(res); a.ldres = Cout
    a.out, a.ldo = out.data_ptr(), Cout
    a.out_nchw_f32 = None
    L.check(L.lib().b2d_conv2d(C.byref(a), L.stream_ptr()), "conv2d")
    torch.cuda.synchronize()
    xin = x if not C2 else torch.cat([x, x2], -1)
    ref = F.conv2d(xin.float().permute(0, 3, 1, 2), w.to(torch.float16).float(), None, stride, k // 2)
    if extras:
        ref = ref + bias[None, :, None, None] + emb[:, :, None, None] + res.float().permute(0, 3, 1, 2)
    ref = ref.permute(0, 2, 3, 1)
    err = (out.float() - ref).norm() / ref.norm()
    print(f"conv N={N} {H}x{W} C=({C1},{C2})->{Cout} k={k} s={stride}: rel err {err.item():.3e}", flush=True)
    return err.item()


def run_wgrad(N, H, W, C1, C2, Cout, k, stride):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(2)
    x = torch.randn(N, H, W, C1, device=dev, generator=g).to(torch.float16)
    x2 = torch.randn(N, H, W, C2, device=dev, generator=g).to(torch.float16) if C2 else None
    Ho, Wo = H // stride, W // stride
    dy = torch.randn(N, Ho, Wo, Cout, device=dev, generator=g).to(torch.float16)
    dw = torch.zeros(Cout, k * k * (C1 + C2), device=dev)
    a = L.WgradArgs()
    a.x, a.ldx, a.C1 = x.data_ptr(), C1, C1
    a.x2, a.ldx2, a.C2 = (x2.data_ptr() if C2 else None), C2, C2
    a.N, a.H, a.W, a.ksize, a.stride = N, H, W, k, stride
    a.dy, a.lddy, a.Cout = dy.data_ptr(), Cout, Cout
    a.dw = dw.data_ptr()
    L.check(L.lib().b2d_conv2d_wgrad(C.byref(a), L.stream_ptr()), "wgrad")
    torch.cuda.synchronize()
    xin = (x if not C2 else torch.cat([x, x2], -1)).float().permute(0, 3, 1, 2).requires_grad_(False)
    w0 = torch.zeros(Cout, C1 + C2, k, k, device=dev, requires_grad=True)
    y = F.conv2d(xin, w0, None, stride, k // 2)
    (gw,) = torch.autograd.grad(y, w0, dy.float().permute(0, 3, 1, 2))
    ref = gw.permute(0, 2, 3, 1).reshape(Cout, -1)
    err = (dw - ref).norm() / ref.norm()
    print(f"wgrad N={N} {H}x{W} C=({C1},{C2})->{Cout} k={k} s={stride}: rel err {err.item():.3e}", flush=True)
    return err.item()


if __name__ == "__main__":
    L.check(L.lib().b2d_check_device(), "device")
    worst = 0.0
    cases = [
        (2, 16, 16, 64, 0, 64, 1, 1),
        (2, 16, 16, 64, 0, 64, 3, 1),
        (4, 8, 8, 128, 0, 128, 3, 1),
        (2, 64, 64, 64, 0, 64, 3, 1),
        (2, 32, 32, 128, 64, 128, 3, 1),
        (2, 32, 32, 64, 0, 64, 3, 2),
        (1, 16, 16, 256, 256, 256, 3, 1),
        (2, 16, 16, 64, 0, 3, 3, 1),
        (2, 16, 16, 128, 0, 384, 1, 1),
        (1, 128, 128, 64, 0, 64, 3, 1),
    ]
    for c in cases:
        worst = max(worst, run_conv(*c))
    for c in cases:
        if c[5] % 64 == 0 and (c[5] == 64 or c[5] % 128 == 0):
            worst = max(worst, run_wgrad(*c))
    print("WORST", worst)
    sys.exit(0 if worst < 1e-2 else 1)
