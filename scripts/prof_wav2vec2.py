"""Developer profile (GPU): the HF wav2vec2-base encoder alone (forward + backward, bf16 autocast, 256 windows of 4000
samples) under torch.profiler -> kernel totals, to see which third-party ops the 13 ms per step go to."""
import collections
import os
import re
import sys
import torch
sys.path.insert(0, ".")
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
from lipreading_video_generation_b200.unet_audio import Wav2Vec2Encoder  # noqa: E402
from torch.profiler import profile, ProfilerActivity  # noqa: E402

torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
enc = Wav2Vec2Encoder().cuda().train()
a = torch.randn(256, 4000, device="cuda")


def run():
    with torch.autocast("cuda", dtype=torch.bfloat16):
        f = enc({"input_values": a}).mean(dim=1)
    f.float().square().sum().backward()


for _ in range(3):
    run()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
    run()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
agg = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    name = re.sub(r"\(.*", "", re.sub(r"<.*", "", e.name))
    agg[name][0] += 1
    agg[name][1] += e.time_range.end - e.time_range.start
tot = sum(v for _, v in agg.values())
print(f"kernel-busy {tot / 1e3:.2f} ms, {sum(n for n, _ in agg.values())} launches")
for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
    print(f"{v / 1e3:8.3f} ms x{n:4d}  {k[:110]}")
print(prof.key_averages(group_by_input_shape=True).table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=40,
                                                          max_shapes_column_width=60))
