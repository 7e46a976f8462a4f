"""Developer check (GPU): attention fwd (and bwd when present) vs fp32 torch on fp16-rounded inputs."""
import math
import sys

import torch

sys.path.insert(0, ".")
from lipreading_video_generation_b200 import _lib as L  # noqa: E402


def ref_attn(qkv, heads, ch, new_order):
    # qkv (N, L, 3C) -> reference layout (N, 3C, L); unet.py:349-366 / 382-401
    N, Lq, W3 = qkv.shape
    x = qkv.float().permute(0, 2, 1)
    if new_order:
        q, k, v = x.chunk(3, dim=1)
        q = q.reshape(N * heads, ch, Lq); k = k.reshape(N * heads, ch, Lq); v = v.reshape(N * heads, ch, Lq)
    else:
        q, k, v = x.reshape(N * heads, ch * 3, Lq).split(ch, dim=1)
    scale = 1 / math.sqrt(math.sqrt(ch))
    w = torch.einsum("bct,bcs->bts", q * scale, k * scale)
    w = torch.softmax(w, dim=-1)
    a = torch.einsum("bts,bcs->bct", w, v)
    return a.reshape(N, heads * ch, Lq).permute(0, 2, 1)  # (N, L, C)


def run(N, Lq, heads, ch, new_order=0, std=1.0, time_it=False):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(3)
    C = heads * ch
    qkv = (torch.randn(N, Lq, 3 * C, device=dev, generator=g) * std).to(torch.float16)
    out = torch.empty(N, Lq, C, device=dev, dtype=torch.float16)
    lse = torch.empty(N, heads, Lq, device=dev)
    lib = L.lib()

    def call():
        L.check(lib.b2d_attention_fwd(qkv.data_ptr(), 3 * C, N, Lq, heads, ch, new_order, out.data_ptr(), C,
                                      lse.data_ptr(), L.stream_ptr()), "attention_fwd")
    call()
    torch.cuda.synchronize()
    ref = ref_attn(qkv, heads, ch, new_order)
    err = ((out.float() - ref).norm() / ref.norm()).item()
    msg = f"attn fwd N={N} L={Lq} heads={heads} ch={ch} new={new_order} std={std}: rel err {err:.3e}"
    if time_it:
        for _ in range(3):
            call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            call()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        fl = 4.0 * Lq * Lq * ch * heads * N
        msg += f"  {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s"
    print(msg, flush=True)
    return err


if __name__ == "__main__":
    L.check(L.lib().b2d_check_device(), "device")
    worst = 0.0
    worst = max(worst, run(2, 256, 1, 64))
    worst = max(worst, run(2, 1024, 1, 128))
    worst = max(worst, run(2, 256, 1, 256))
    worst = max(worst, run(1, 4096, 1, 64, std=2.0))
    worst = max(worst, run(2, 320, 2, 64))          # ragged L, 2 heads legacy order
    worst = max(worst, run(2, 200, 2, 64, 1))       # ragged L, new order
    worst = max(worst, run(1, 64, 1, 256))          # L smaller than a tile
    worst = max(worst, run(1, 1024, 1, 64, std=4.0))  # large logits: exercises the lazy rescale
    run(64, 4096, 1, 64, time_it=True)
    run(256, 4096, 1, 64, time_it=True)        # bench shape (cfg2 full-res layer)
    run(256, 1024, 1, 128, time_it=True)       # cfg2 half-res layer
    run(64, 1024, 1, 128, time_it=True)
    run(64, 256, 1, 256, time_it=True)
    run(8, 16384, 1, 64, time_it=True)
    print("WORST", worst)
    sys.exit(0 if worst < 1e-2 else 1)
