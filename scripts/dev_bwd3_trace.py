"""Developer timeline of the fused attention backward (CTA (0,0,0)): build csrc/attention_bwd3.cu with -DB3_TRACE
(B2D_EXTRA_NVCC_FLAGS=-DB3_TRACE), then  python scripts/dev_bwd3_trace.py run gpurun_out/bwd3_trace.bin  on the GPU box
(the script owns the device buffer and hands its address to the library through B2D_BWD3_TRACE_PTR) and
python scripts/dev_bwd3_trace.py show gpurun_out/bwd3_trace.bin  anywhere."""
import struct
import sys


def run(path):
    import os
    import torch
    sys.path.insert(0, ".")
    buf = torch.zeros(1024, dtype=torch.int64, device="cuda")
    os.environ["B2D_BWD3_TRACE_PTR"] = str(buf.data_ptr())
    from lipreading_video_generation_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(0)
    N, H, C = 32, 64, 64
    qkv = torch.randn(N, H, H, 3 * C, device="cuda", generator=g).half()
    d_o = torch.randn(N, H, H, C, device="cuda", generator=g).half()
    out, lse = ops.attention_fwd(qkv, 1)
    for _ in range(2):
        ops.attention_bwd(qkv, out, d_o, lse, 1)
    torch.cuda.synchronize()
    open(path, "wb").write(struct.pack("1024q", *buf.cpu().tolist()))


def show(path):
    v = struct.unpack("1024q", open(path, "rb").read())
    t0 = min(x for x in v if x > 0)
    names = ["wait S", "ldtm", "math", "wait pair", "pack+st+fence", "stage dQ", "issue dQ"]
    for b in range(2):
        for m in list(range(1, 6)) + list(range(14, 18)):
            row = [v[(b * 32 + m) * 8 + k] - t0 for k in range(8)]
            d = [row[i + 1] - row[i] for i in range(7)]
            print(f"group {b} step {m:2d} start {row[0]:7d} | " + " ".join(f"{n} {x:5d}" for n, x in zip(names, d)) +
                  f" | total {row[7] - row[0]:5d}")
    names = ["wait p_ready", "issue dV,dK", "issue dQ", "issue S,dP"]
    for j in list(range(2, 10)) + list(range(28, 34)):
        row = [v[512 + j * 6 + k] - t0 for k in range(5)]
        print(f"issuer j {j:2d} start {row[0]:7d} | " + " ".join(f"{n} {row[i + 1] - row[i]:5d}" for i, n in enumerate(names)))


if __name__ == "__main__":
    run(sys.argv[2]) if sys.argv[1] == "run" else show(sys.argv[2])
