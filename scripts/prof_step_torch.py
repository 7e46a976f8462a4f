"""Developer profile (GPU): one train64 step under torch.profiler (CUPTI activity records: kernel durations on the
live timeline, not ncu's serialised cold-cache replays) -> per-kernel totals and the GPU idle time inside the step."""
import os
import re
import sys
import collections
import torch
sys.path.insert(0, ".")
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
import bench  # noqa: E402

torch.backends.cudnn.benchmark = True
dev = torch.device("cuda:0")
S, n_img = 64, 256
model = bench.build_model(S, dev)
step = bench.TrainStep(model, S, 1, dev)
x0, cond, audio = bench.make_batch(n_img, S, 1234, dev)
t = torch.randint(0, 1000, (n_img,)).to(dev)
for _ in range(4):
    step(x0, cond, audio, t)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        step(x0, cond, audio, t)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
agg = collections.defaultdict(lambda: [0, 0.0])
busy, t_end = 0.0, None
first, last = ev[0].time_range.start, max(e.time_range.end for e in ev)
for e in ev:
    d = e.time_range.end - e.time_range.start
    name = re.sub(r"<.*", "", e.name)
    name = re.sub(r"\(.*", "", name)
    agg[name][0] += 1
    agg[name][1] += d
    busy += d
print(f"span {(last - first) / 2e3:.2f} ms/step   kernel-busy {busy / 2e3:.2f} ms/step   idle {(last - first - busy) / 2e3:.2f} ms/step")
for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:40]:
    print(f"{v / 2e3:8.3f} ms/step x{n // 2:4d}  {k[:100]}")
