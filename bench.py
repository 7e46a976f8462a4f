"""bench.py -- denoiser train step throughput (clips/s), BASELINE.json configs[1]:
single B200, batch 16 clips x 16 frames 64x64 (N = 256 images folded into the batch), bf16 train step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload train64|train128|sample128]

One "step" = q_sample + UNetAudio forward (wav2vec2 + sm_100a torso) + MSE + backward + Adam step on one
batch of synthetic clips (train.py:109-134).  Prints ONE JSON line (see DESIGN.md "measurement").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")   # no network: wav2vec2-base weights are random-init
os.environ.setdefault("NCCL_DEBUG", "WARN")              # keep stdout to the one JSON line

import torch  # noqa: E402

WORKLOADS = {
    # name: (clips per GPU, frames per clip, S, kind)
    "train64": (16, 16, 64, "train"),      # BASELINE configs[1]
    "train128": (16, 16, 128, "train"),    # BASELINE configs[2] per-GPU shape
    "sample128": (8, 16, 128, "sample"),   # BASELINE configs[3] per-GPU shape
    "train64_1clip": (1, 16, 64, "train"),  # diagnostic: 16x less GPU work, same launches -> host cost of one step
    "train3d_cfg5": (1, 64, 128, "train3d"),   # BASELINE configs[4]: UNetModel(dims=3) on one (195, 64, 128, 128) clip
    "train3d_small": (1, 16, 64, "train3d"),   # the same path at (195, 16, 64, 64)
}
FWD_GFLOP_PER_IMAGE = {64: 43.87, 128: 470.21}     # SURVEY section 8(d), algorithmic, UNet torso


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured"
    return 1590.0, 1400.0, 6650.0, "fallback"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index):
        self.index, self.rows, self._stop, self._th = index, [], threading.Event(), None

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self._stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([c.strip() for c in o.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._th.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def make_batch(n_img, S, seed, device, pinned=False):
    """SURVEY section 8(d) synthetic inputs: x0, reference image ~ U(-1,1); audio ~ N(0,1) (N,4000); one clip shares
    its frame-0 reference image across its frames."""
    g = torch.Generator().manual_seed(seed)
    x0 = torch.rand(n_img, 3, S, S, generator=g) * 2 - 1
    cond = torch.rand(n_img, 3, S, S, generator=g) * 2 - 1
    audio = torch.randn(n_img, 4000, generator=g)
    if pinned:
        return x0.pin_memory(), cond.pin_memory(), audio.pin_memory()
    return x0.to(device), cond.to(device), audio.to(device)


def build_model(S, device, seed=0):
    import lipreading_video_generation_b200 as P
    torch.manual_seed(seed)
    model = P.UNetAudio(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
                        attention_resolutions=(1, 2, 4), audio_feature_dim=768, projected_audio_dim=128)   # train.py:88-97
    # un_zero: zero_module'd tensors get N(0, 0.02) so that no layer's work is skipped by zeros
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if (".out_layers.3." in name) or (".proj_out." in name) or name.startswith("out.2."):
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    return model.to(device)


class SampleStep:
    """test.py:57-65 on the product: one reverse-diffusion step for a batch of frames -- model forward (no autograd
    graph, conditioning encoded once and reused: it is loop invariant) + DDPM posterior update with fresh noise."""

    def __init__(self, model, S, device):
        import lipreading_video_generation_b200 as P
        self.model, self.device = model.eval(), device
        self.sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
        self.i = 999
        self.feats = None

    def prepare(self, audio):
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            self.feats = self.model.encode_audio({"input_values": audio}).float()

    @torch.no_grad()
    def __call__(self, xt, cond):
        i = self.i
        self.i = i - 1 if i > 1 else 999
        t = torch.full((xt.shape[0],), i, device=self.device, dtype=torch.long)
        eps = self.model(xt, cond, None, t, audio_features=self.feats)
        x_prev, x0 = self.sched.sample_prev_timestep(xt, eps, i)
        return x_prev, x0


class TrainStep:
    """train.py:109-134 on the product: zero_grad, q_sample, model, MSE, backward, (grad all-reduce), Adam."""

    def __init__(self, model, S, world, device):
        import lipreading_video_generation_b200 as P
        from lipreading_video_generation_b200.parallel import GradientAllReducer
        self.P, self.model, self.device, self.world = P, model, device, world
        self.sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
        self.crit = P.MSELoss()
        if os.environ.get("B2D_TORCH_ADAM") == "1":          # A/B runs only
            self.opt = torch.optim.Adam(model.parameters(), lr=1e-2, fused=True)
        else:
            self.opt = P.Adam(model.parameters(), lr=1e-2)                            # train.py:102, one launch
        self.reducer = GradientAllReducer(model) if world > 1 else None
        model.train()

    def __call__(self, x0, cond, audio, t):
        self.opt.zero_grad(set_to_none=True)
        noise = torch.randn_like(x0)                                                  # train.py:122
        xt = self.sched.add_noise(x0, noise, t)
        with torch.autocast("cuda", dtype=torch.bfloat16):                           # HF wav2vec2 only; torso is ours
            feats = self.model.encode_audio({"input_values": audio})
        pred = self.model(xt, cond, None, t, audio_features=feats.float())
        loss = self.crit(pred, noise)
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
        self.opt.step()
        return loss


def cpu_baseline(S, threads, seconds_budget=25.0):
    """The oracle port (CPU restatement of the reference, fp32) timed on the host: fwd + bwd of the torso on a
    bounded sample of the same workload (images of the same shape, as many as fit the time budget)."""
    from oracle import restatement as R
    import lipreading_video_generation_b200 as P
    torch.set_num_threads(threads)

    class Stub(torch.nn.Module):
        def forward(self, a):
            return a["feats"]
    m = P.UNetAudio(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
                    attention_resolutions=(1, 2, 4), dropout=0.0, audio_feature_dim=768, projected_audio_dim=128,
                    audio_encoder=Stub())
    sd = R.protocol_params(m.state_dict(), seed=0)
    cfg = dict(image_size=S, model_channels=64, num_res_blocks=2, attention_resolutions=(1, 2, 4), channel_mult=(1, 2, 4), num_heads=1)
    n = 2
    g = torch.Generator().manual_seed(0)
    xt = torch.randn(n, 3, S, S, generator=g); cond = torch.rand(n, 3, S, S, generator=g)
    fm = torch.randn(n, 768, generator=g); eps = torch.randn(n, 3, S, S, generator=g); t = torch.randint(0, 1000, (n,), generator=g)

    def one():
        sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        loss = R.mse(R.unet_audio_forward(sdr, cfg, xt, cond, fm, t), eps)
        loss.backward()
    one()                                   # warm-up
    times = []
    t_end = time.time() + seconds_budget
    while len(times) < 5 and time.time() < t_end:
        t0 = time.time(); one(); times.append(time.time() - t0)
    med = sorted(times)[len(times) // 2]
    return n, med, len(times)


def main_sample(a, model, S, clips, frames, world, rank, local, device, config, cores):
    """sampler frames/s (BASELINE configs[3] per-GPU shape): replicas only, no collective on the data path."""
    from lipreading_video_generation_b200 import _lib as L, ops
    n_img = clips * frames
    step = SampleStep(model, S, device)
    x0, cond, audio = make_batch(n_img, S, 1234 + rank, device)
    hx0, hcond, haudio = make_batch(n_img, S, 1234 + rank, device, pinned=True)
    step.prepare(audio)
    xt = torch.randn(n_img, 3, S, S, device=device)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
    for _ in range(a.warmup):
        xt, _ = step(xt, cond)
    barrier()
    l0 = L.lib().b2d_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk, ops.KernelTimer({"attention_fwd", "conv2d", "groupnorm_fwd"}) as kt:
        e0.record()
        for _ in range(a.steps):
            xt, x0p = step(xt, cond)
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1) / a.steps
    launches = (L.lib().b2d_launch_count() - l0) // a.steps
    # end to end: conditioning + x_T from pinned host memory, wav2vec2 encoding, K steps, x0_pred back to the host
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    hxt = torch.randn(n_img, 3, S, S).pin_memory()
    hout = torch.empty(n_img, 3, S, S).pin_memory()
    f0.record()
    dc = hcond.to(device, non_blocking=True); da = haudio.to(device, non_blocking=True); dx = hxt.to(device, non_blocking=True)
    step.prepare(da)
    for _ in range(a.steps):
        dx, x0p = step(dx, dc)
    hout.copy_(x0p, non_blocking=True)
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1) / a.steps
    if world > 1:
        tt = torch.tensor([ms, ms_e2e], device=device)
        torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        ms, ms_e2e = tt.tolist()
        torch.distributed.destroy_process_group()
    if rank != 0:
        return
    burst, sustained, hbm, which = peaks()
    tot = kt.totals()
    by_class, dom = {}, None
    for (name, meta), (t_ms, cnt) in tot.items():
        by_class[name] = by_class.get(name, 0.0) + t_ms / a.steps
        if name == "attention_fwd":
            Nn, Lq, heads, ch = meta
            rec = dict(kernel=f"{name} L={Lq} ch={ch}", ms_per_launch=t_ms / cnt, tflops=4.0 * Lq * Lq * ch * heads * Nn / (t_ms / cnt) / 1e9,
                       ms_per_step=t_ms / a.steps)
            if dom is None or rec["ms_per_step"] > dom["ms_per_step"]:
                dom = rec
    T = 1000
    value = n_img * world / (T * ms / 1000.0)
    fwd_tflop = FWD_GFLOP_PER_IMAGE[S] * n_img / 1000.0
    out = {"metric": "sampler_frames_per_s", "value": round(value, 4), "unit": "frames/s (1000-step DDPM)", "n_gpus": world,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": config,
           "frames_per_s_50_step_ddim": round(n_img * world / (50 * ms / 1000.0), 3),
           "e2e": {"value": round(n_img * world / (T * ms_e2e / 1000.0), 4), "unit": "frames/s (1000-step DDPM)",
                   "h2d_bytes_per_step": sum(t.numel() * t.element_size() for t in (hcond, haudio, hxt)) // a.steps,
                   "d2h_bytes_per_step": hout.numel() * 4 // a.steps, "ms_per_step": round(ms_e2e, 3)},
           "gpu_launches": int(launches),
           "roofline": {"bound": "tensor", "achieved": round(dom["tflops"], 1), "peak": sustained, "unit": "TFLOP/s",
                        "frac": round(dom["tflops"] / sustained, 4), "traffic": None, "kernel": dom["kernel"],
                        "ms_per_launch": round(dom["ms_per_launch"], 4), "share_of_step": round(dom["ms_per_step"] / ms, 3),
                        "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({which}); algorithmic FLOPs 4*L^2*ch*N"},
           "step_tflops_algorithmic": round(fwd_tflop / (ms / 1000.0), 1),
           "step_frac_of_sustained_peak": round(fwd_tflop / (ms / 1000.0) / sustained, 4),
           "kernel_class_ms_per_step": {k: round(v, 3) for k, v in sorted(by_class.items())},
           "cpu_baseline": None, "clocks": clk.summary()}
    print(json.dumps(out))


def main_train3d(a, S, clips, frames, world, rank, local, device, config, cores):
    """UNetModel(dims=3) train step on whole clips: x (B, 195, T, S, S) -> eps (B, 3, T, S, S),
    MSE, backward, Adam.  Attention runs over L = T*S*S positions per clip (1,048,576 at cfg5), which the reference
    cannot materialise.  With N > 1 GPUs the ONE clip is split across the ranks (row N1, strong scaling): attention is
    sharded by query rows (parallel.SequenceParallelAttention: all-gather of the output, all-reduce of dqkv), the rest of
    the torso runs replicated on identical inputs, parameter gradients are averaged to keep the replicas identical."""
    import lipreading_video_generation_b200 as P
    from lipreading_video_generation_b200 import _lib as L, ops
    torch.manual_seed(0)
    model = P.UNetModel(image_size=S, in_channels=195, model_channels=64, out_channels=3, num_res_blocks=2,
                        attention_resolutions=(1, 2, 4), dropout=0.1, channel_mult=(1, 2, 4), dims=3)
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if (".out_layers.3." in name) or (".proj_out." in name) or name.startswith("out.2."):
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    model = model.to(device).train()
    opt = P.Adam(model.parameters(), lr=1e-2)
    crit = P.MSELoss()
    reducer = sp = None
    if world > 1:
        from lipreading_video_generation_b200.parallel import (GradientAllReducer, SequenceParallelAttention,
                                                               broadcast_parameters)
        broadcast_parameters(model)
        sp = SequenceParallelAttention(model)
        reducer = GradientAllReducer(model)
        config["parallelism"] = f"sp{world}: one clip, attention sharded by query rows, torso replicated"
    gi = torch.Generator().manual_seed(1234)              # every rank works on the SAME clip
    hx = torch.randn(clips, 195, frames, S, S, generator=gi).pin_memory()
    heps = torch.randn(clips, 3, frames, S, S, generator=gi).pin_memory()
    ht = torch.randint(0, 1000, (clips,), generator=gi).pin_memory()
    x, eps, t = hx.to(device), heps.to(device), ht.to(device)

    def step(x_, eps_, t_):
        opt.zero_grad(set_to_none=True)
        loss = crit(model(x_, t_), eps_)
        loss.backward()
        if reducer is not None:
            reducer.finish()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        step(x, eps, t)
    barrier()
    l0 = L.lib().b2d_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk, ops.KernelTimer({"attention_fwd", "attention_bwd"}) as kt:
        e0.record()
        for _ in range(a.steps):
            step(x, eps, t)
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1) / a.steps
    launches = (L.lib().b2d_launch_count() - l0) // a.steps
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(a.steps):
        loss = step(hx.to(device, non_blocking=True), heps.to(device, non_blocking=True), ht.to(device, non_blocking=True))
        loss_host = loss.item()
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1) / a.steps
    with ops.KernelTimer({"attention_fwd", "attention_bwd", "conv2d", "conv2d_wgrad", "groupnorm_fwd", "groupnorm_bwd"}) as kt_all:
        step(x, eps, t)
        barrier()
    if world > 1:
        tt = torch.tensor([ms, ms_e2e], device=device)
        torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        ms, ms_e2e = tt.tolist()
        torch.distributed.destroy_process_group()
    if rank != 0:
        return
    burst, sustained, hbm, which = peaks()
    dom, by_class = None, {}
    for (name, meta), (t_ms, cnt) in kt_all.totals().items():
        by_class[name] = by_class.get(name, 0.0) + t_ms
    for (name, meta), (t_ms, cnt) in kt.totals().items():
        Nn, Lk, heads, ch = meta[:4]
        Lq = meta[4] if len(meta) > 4 else Lk             # a sequence-parallel shard visits Lq of the Lk query rows
        flops = (4.0 if name == "attention_fwd" else 10.0) * Lq * Lk * ch * heads * Nn
        rec = dict(kernel=f"{name} L={Lk} ch={ch}" + (f" (query rows {Lq} of {Lk} per GPU)" if Lq != Lk else ""),
                   ms_per_launch=t_ms / cnt, tflops=flops / (t_ms / cnt) / 1e9, ms_per_step=t_ms / a.steps)
        if dom is None or rec["ms_per_step"] > dom["ms_per_step"]:
            dom = rec
    h2d = sum(v.numel() * v.element_size() for v in (hx, heps, ht))
    out = {"metric": "denoiser_train_clips_per_s", "value": round(clips / (ms / 1000.0), 4), "unit": "clips/s",
           "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
           "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": config,
           "seq_parallel_bytes_per_step": (sp.bytes_moved // (a.warmup + 2 * a.steps + 1)) if sp is not None else 0,
           "e2e": {"value": round(clips / (ms_e2e / 1000.0), 4), "unit": "clips/s", "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e, 3)},
           "gpu_launches": int(launches),
           "roofline": {"bound": "tensor", "achieved": round(dom["tflops"], 1), "peak": sustained, "unit": "TFLOP/s",
                        "frac": round(dom["tflops"] / sustained, 4), "traffic": None, "kernel": dom["kernel"],
                        "ms_per_launch": round(dom["ms_per_launch"], 4), "share_of_step": round(dom["ms_per_step"] / ms, 3),
                        "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({which}); algorithmic FLOPs 4*L^2*ch fwd, 10*L^2*ch bwd"},
           "kernel_class_ms_per_step": {k: round(v, 3) for k, v in sorted(by_class.items())},
           "cpu_baseline": None, "clocks": clk.summary(), "last_loss": loss_host}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="train64", choices=list(WORKLOADS))
    a = ap.parse_args()
    clips, frames, S, kind = WORKLOADS[a.workload]
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    config = {"workload": f"{a.workload}: denoiser {kind} step, {clips} clips x {frames} frames {S}x{S} per GPU "
                          f"(N={clips * frames} images folded into the batch), UNetAudio train.py config + wav2vec2-base",
              "clips_per_gpu": clips, "frames": frames, "image": S, "parallelism": f"dp{world}",
              "l2_policy": "inputs+activations >> L2 (126 MB): >10 GB touched per step"}
    if kind == "train3d":
        config["workload"] = (f"{a.workload}: UNetModel(dims=3) train step on {clips} clip(s) (195, {frames}, {S}, {S}) per GPU, "
                              f"train.py torso config (64 ch, mult (1,2,4), 2 res blocks, attention at all levels over "
                              f"L = T*S*S = {frames * S * S} positions), MSE + backward + Adam, no audio encoder")

    if a.impl == "reference":
        # The reference is PyTorch-CPU here (/root/reference does not travel): its oracle port, all host threads.
        if rank != 0:
            return
        vals = []
        for _ in range(max(1, min(a.steps, 3))):
            n, sec, reps = cpu_baseline(S, cores, seconds_budget=40.0)
            vals.append((n / frames) / sec)
        v = sorted(vals)[len(vals) // 2]
        print(json.dumps({"impl": "reference", "metric": "denoiser_train_clips_per_s", "value": v, "unit": "clips/s",
                          "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1000.0 * (2 / frames) / v,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": "clips/s", "cores": cores, "kind": "port",
                                           "sample": f"2 images {S}x{S} (1/{frames // 2 if frames >= 2 else 1} clip): torso fwd+bwd fp32, no wav2vec2/Adam"},
                          "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    assert torch.cuda.is_available(), "bench.py needs a B200; there is no CPU fallback for the product path"
    torch.backends.cudnn.benchmark = True      # third-party wav2vec2 (HF / cuDNN) only: let cuDNN pick its conv kernels
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    from lipreading_video_generation_b200 import _lib as L, ops
    from lipreading_video_generation_b200.parallel import broadcast_parameters
    L.check(L.lib().b2d_check_device(), "b2d_check_device")
    n_img = clips * frames
    if kind == "train3d":
        return main_train3d(a, S, clips, frames, world, rank, local, device, config, cores)
    model = build_model(S, device)
    if world > 1:
        broadcast_parameters(model)
    if kind == "sample":
        return main_sample(a, model, S, clips, frames, world, rank, local, device, config, cores)
    step = TrainStep(model, S, world, device)
    x0, cond, audio = make_batch(n_img, S, 1234 + rank, device)
    hx0, hcond, haudio = make_batch(n_img, S, 1234 + rank, device, pinned=True)
    gt = torch.Generator().manual_seed(99 + rank)
    t_dev = torch.randint(0, 1000, (n_img,), generator=gt).to(device)
    t_host = torch.randint(0, 1000, (n_img,), generator=gt).pin_memory()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        step(x0, cond, audio, t_dev)
    # ---------------- timed region 1: inputs resident in HBM ("value") + roofline events on attention kernels
    barrier()
    l0 = L.lib().b2d_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # only the candidate dominant kernels carry events inside the timed region (26 launches per step); the full
    # per-class breakdown (~530 event pairs per step, which costs 2-3 ms of step time) runs in an untimed pass below
    with ClockSampler(local) as clk, ops.KernelTimer({"attention_fwd", "attention_bwd"}) as kt:
        e0.record()
        h0 = time.perf_counter()
        for _ in range(a.steps):
            step(x0, cond, audio, t_dev)
        host_ms = (time.perf_counter() - h0) * 1e3 / a.steps     # host time to enqueue one step (no sync inside)
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1) / a.steps
    launches = (L.lib().b2d_launch_count() - l0) // a.steps
    # ---------------- timed region 2: end to end through the public API with HOST buffers
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    # the loss of every step is copied to pinned host memory and READ on the host one step later (after the next step
    # has been enqueued), the last one before the timer stops: logging the loss does not drain the launch queue
    loss_pin = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    losses_host = []
    for i in range(a.steps):
        dx0 = hx0.to(device, non_blocking=True); dc = hcond.to(device, non_blocking=True)
        da = haudio.to(device, non_blocking=True); dt = t_host.to(device, non_blocking=True)
        loss = step(dx0, dc, da, dt)
        loss_pin[i % 2].copy_(loss.detach(), non_blocking=True)     # device -> host read of the step's result
        loss_ev[i % 2].record()
        if i > 0:
            loss_ev[(i - 1) % 2].synchronize()
            losses_host.append(float(loss_pin[(i - 1) % 2]))
    loss_ev[(a.steps - 1) % 2].synchronize()
    losses_host.append(float(loss_pin[(a.steps - 1) % 2]))
    loss_host = losses_host[-1]
    assert len(losses_host) == a.steps
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1) / a.steps
    # ---------------- untimed diagnostic pass: per-class kernel time of one step
    with ops.KernelTimer({"attention_fwd", "attention_bwd", "conv2d", "conv2d_wgrad", "groupnorm_fwd", "groupnorm_bwd"}) as kt_all:
        step(x0, cond, audio, t_dev)
        barrier()
    if world > 1:
        tt = torch.tensor([ms, ms_e2e], device=device)
        torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        ms, ms_e2e = tt.tolist()
    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return
    burst, sustained, hbm, which = peaks()
    # roofline of the dominant kernel class: full-resolution (ch = 64) attention, forward
    tot = kt.totals()
    dom = None
    by_class = {}
    for (name, meta), (t_ms, cnt) in kt_all.totals().items():
        by_class[name] = by_class.get(name, 0.0) + t_ms
        print(f"[kernel] {name:14s} {str(meta):40s} x{cnt:3d}/step  {t_ms / cnt:8.4f} ms/launch  {t_ms:8.3f} ms/step",
              file=sys.stderr)
    for (name, meta), (t_ms, cnt) in tot.items():
        if not name.startswith("attention"):
            continue
        Nn, Lq, heads, ch = meta
        flops = (4.0 if name == "attention_fwd" else 10.0) * Lq * Lq * ch * heads * Nn
        rec = dict(kernel=f"{name} L={Lq} ch={ch}", ms_per_launch=t_ms / cnt, launches_per_step=cnt // a.steps,
                   tflops=flops / (t_ms / cnt) / 1e9, ms_per_step=t_ms / a.steps)
        if dom is None or rec["ms_per_step"] > dom["ms_per_step"]:
            dom = rec
    traffic = None          # DRAM bytes per launch of the dominant kernel from one `ncu --set full` capture
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_traffic.json")) as f:
            tj = json.load(f)
        if dom["kernel"] == "attention_bwd L=4096 ch=64" and n_img == 256:
            traffic = tj["traffic_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        pass
    roof = {"bound": "tensor", "achieved": round(dom["tflops"], 1), "peak": sustained, "unit": "TFLOP/s",
            "frac": round(dom["tflops"] / sustained, 4), "traffic": traffic, "kernel": dom["kernel"],
            "ms_per_launch": round(dom["ms_per_launch"], 4), "share_of_step": round(dom["ms_per_step"] / ms, 3),
            "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({which}); algorithmic FLOPs 4*L^2*ch*N fwd, 10*L^2*ch*N bwd"}
    if os.environ.get("B2D_BENCH_NO_CPU") == "1":          # profiling runs only (ncu): skip the host-side leg
        n_cpu, sec_cpu, reps = 2, float("inf"), 0
    else:
        n_cpu, sec_cpu, reps = cpu_baseline(S, cores)
    cpu_v = (n_cpu / frames) / sec_cpu
    value = clips * world / (ms / 1000.0)
    train_tflop = 3 * FWD_GFLOP_PER_IMAGE[S] * n_img / 1000.0
    h2d = sum(t.numel() * t.element_size() for t in (hx0, hcond, haudio, t_host))
    out = {"metric": "denoiser_train_clips_per_s", "value": round(value, 3), "unit": "clips/s", "n_gpus": world,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": config,
           "e2e": {"value": round(clips * world / (ms_e2e / 1000.0), 3), "unit": "clips/s", "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e, 3)},
           "gpu_launches": int(launches), "host_enqueue_ms_per_step": round(host_ms, 3), "roofline": roof,
           "step_tflops_algorithmic": round(train_tflop / (ms / 1000.0), 1),
           "step_frac_of_sustained_peak": round(train_tflop / (ms / 1000.0) / sustained, 4),
           "cpu_baseline": {"value": round(cpu_v, 5), "unit": "clips/s", "cores": cores, "kind": "port",
                            "sample": f"{n_cpu} images {S}x{S}: oracle torso fwd+bwd fp32, median of {reps}, {sec_cpu:.2f} s"},
           "kernel_class_ms_per_step": {k: round(v, 3) for k, v in sorted(by_class.items())},
           "clocks": clk.summary(), "last_loss": loss_host}
    print(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
