"""Host cost of each phase of one train step.  With 1 clip the GPU work is small, so the synchronised wall time of a
phase is close to its host enqueue time; with 16 clips the same phases are GPU-bound.
usage: python scripts/dev_host_phases.py [clips]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
import torch
import bench
import lipreading_video_generation_b200 as P

clips = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda:0")
model = bench.build_model(64, dev)
step = bench.TrainStep(model, 64, 1, dev)
n = clips * 16
x0, cond, audio = bench.make_batch(n, 64, 1234, dev)
t = torch.randint(0, 1000, (n,), device=dev)
for _ in range(3):
    step(x0, cond, audio, t)
torch.cuda.synchronize()
acc = {}
def tick(name, t0, sync):
    if sync:
        torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    return time.perf_counter()
for sync in (True, False):
    acc.clear()
    R = 5
    torch.cuda.synchronize()
    w0 = time.perf_counter()
    for _ in range(R):
        t0 = time.perf_counter()
        step.opt.zero_grad(set_to_none=True)
        noise = torch.randn_like(x0)
        xt = step.sched.add_noise(x0, noise, t)
        t0 = tick("q_sample", t0, sync)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            feats = model.encode_audio({"input_values": audio})
        t0 = tick("wav2vec2_fwd", t0, sync)
        pred = model(xt, cond, None, t, audio_features=feats.float())
        loss = step.crit(pred, noise)
        t0 = tick("torso_fwd", t0, sync)
        loss.backward()
        t0 = tick("backward", t0, sync)
        step.opt.step()
        t0 = tick("adam", t0, sync)
    torch.cuda.synchronize()
    tot = (time.perf_counter() - w0) * 1e3 / R
    print(f"clips={clips} sync_per_phase={sync} total {tot:.2f} ms/step: " + ", ".join(f"{k} {v / R:.2f}" for k, v in acc.items()))
