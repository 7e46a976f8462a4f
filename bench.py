"""bench.py -- BASELINE.json's metric: denoiser train clips/s and sampler frames/s on N B200s, % of roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
                    [--workload train64|train128|sample128|sample128_b1|train64_1clip|train3d_cfg5|train3d_small] [--no-extras]

The default line is `train64` (BASELINE configs[1]: 16 clips x 16 frames 64x64 per GPU, N = 256 images folded into the
batch); the SAME run then measures, as `extra` sub-records, the other half of the metric -- `sample128` (configs[3]:
1000-step DDPM and a real 50-step DDIM loop, frames/s) -- and `train128` (configs[2] per-GPU shape), plus the achieved
HBM rate of the elementwise kernels (q_sample, MSE, sampler update) at the configs[2]/[3] sizes.

One train "step" = q_sample + UNetAudio forward (wav2vec2 + sm_100a torso) + MSE + backward + gradient all-reduce + Adam
on one batch of synthetic clips (train.py:109-134); one sampler "step" = model forward + sample_prev_timestep
(test.py:57-65).  Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")   # no network: wav2vec2-base weights are random-init

import torch  # noqa: E402

WORKLOADS = {
    # name: (clips per GPU, frames per clip, S, kind)
    "train64": (16, 16, 64, "train"),      # BASELINE configs[1]
    "train128": (16, 16, 128, "train"),    # BASELINE configs[2] per-GPU shape
    "sample128": (8, 16, 128, "sample"),   # BASELINE configs[3] per-GPU shape
    "sample128_b1": (1, 1, 128, "sample"),  # the reference's own sampling operating point (test.py:54): batch 1
    "train64_1clip": (1, 16, 64, "train"),  # diagnostic: 16x less GPU work, same launches -> host cost of one step
    "train3d_cfg5": (1, 64, 128, "train3d"),   # BASELINE configs[4]: UNetModel(dims=3) on one (195, 64, 128, 128) clip
    "train3d_small": (1, 16, 64, "train3d"),   # the same path at (195, 16, 64, 64)
}
FWD_GFLOP_PER_IMAGE = {64: 43.87, 128: 470.21}     # SURVEY section 8(d), algorithmic, UNet torso
CLASSES = {"attention_fwd", "attention_bwd", "conv2d", "conv2d_wgrad", "groupnorm_fwd", "groupnorm_bwd"}


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
    graph, conditioning encoded once and reused: it is loop invariant) + the sampler update with fresh noise.
    mode "ddpm": LinearNoiseScheduler.sample_prev_timestep over t = 999..0; mode "ddim": the 50-step eta = 0
    subsequence of the same schedule through scheduler.ddim_step."""

    def __init__(self, model, S, device, mode="ddpm", ddim_steps=50, graph=None):
        import lipreading_video_generation_b200 as P
        self.model, self.device, self.mode = model.eval(), device, mode
        self.sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
        self.seq = [int(round(v)) for v in torch.linspace(999, 0, ddim_steps).tolist()]
        self.k = 0
        self.feats = None
        # the model forward as one CUDA graph launch (sampling.GraphedDenoiser); B2D_SAMPLE_GRAPH=0 = eager launches
        self.use_graph = (os.environ.get("B2D_SAMPLE_GRAPH", "1") != "0") if graph is None else graph
        self.graphed = None

    def prepare(self, audio):
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            self.feats = self.model.encode_audio({"input_values": audio}).float()

    def _eps(self, xt, cond, i):
        if self.use_graph:
            import lipreading_video_generation_b200 as P
            if self.graphed is None:
                self.graphed = P.GraphedDenoiser(self.model, cond, self.feats)
                self._cond_feats = (cond, self.feats)
            elif self._cond_feats[0] is not cond or self._cond_feats[1] is not self.feats:
                self.graphed.set_conditioning(cond, self.feats)
                self._cond_feats = (cond, self.feats)
            return self.graphed(xt, i)
        t = torch.full((xt.shape[0],), i, device=self.device, dtype=torch.long)
        return self.model(xt, cond, None, t, audio_features=self.feats)

    @torch.no_grad()
    def __call__(self, xt, cond):
        if self.mode == "ddpm":
            i = 999 - (self.k % 999)                # never t = 0 inside a timed region: every step draws noise
            self.k += 1
            return self.sched.sample_prev_timestep(xt, self._eps(xt, cond, i), i)
        k = self.k % len(self.seq)
        self.k += 1
        i, prev = self.seq[k], (self.seq[k + 1] if k + 1 < len(self.seq) else -1)
        return self.sched.ddim_step(xt, self._eps(xt, cond, i), i, prev)


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
            self.opt = P.Adam(model.parameters(), lr=1e-2, skip_on_overflow=model)     # train.py:102, one launch
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

    def close(self):
        if self.reducer is not None:
            self.reducer.detach()


# ----------------------------------------------------------------------------- CPU baselines (oracle port, host cores)
def _cpu_name():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


class CpuTrainCfg1:
    """BASELINE.md section 2 / SURVEY 8(d) 'config 1 exactly': the reference's train step (train.py:109-134) on the host
    cores -- 2 clips x 8 frames = 16 images 64x64, fp32, HF wav2vec2-base (random init, train mode) + the oracle port of
    the UNet torso (AttentionBlock under its always-on activation checkpoint, unet.py:309) + MSE + backward + Adam
    over all 110.7 M parameters.  /root/reference does not exist on the GPU box, so the torso is oracle/restatement.py
    (pinned against the real reference by oracle/make_golden*.py): kind = "port"."""
    N, S = 16, 64

    def __init__(self, threads):
        from transformers import Wav2Vec2Config, Wav2Vec2Model
        from oracle import restatement as R
        import lipreading_video_generation_b200 as P
        torch.set_num_threads(threads)
        self.R = R

        class Stub(torch.nn.Module):
            def forward(self, a):
                return a["feats"]
        shell = P.UNetAudio(image_size=self.S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
                            attention_resolutions=(1, 2, 4), audio_feature_dim=768, projected_audio_dim=128,
                            audio_encoder=Stub())
        sd = R.protocol_params(shell.state_dict(), seed=0)
        self.sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        torch.manual_seed(0)
        self.enc = Wav2Vec2Model(Wav2Vec2Config()).train()                     # unet_audio.py:14 (random init: no network)
        self.cfg = dict(image_size=self.S, model_channels=64, num_res_blocks=2, attention_resolutions=(1, 2, 4),
                        channel_mult=(1, 2, 4), num_heads=1, checkpoint_attention=True)
        self.opt = torch.optim.Adam(list(self.sd.values()) + list(self.enc.parameters()), lr=1e-2)   # train.py:102
        self.tab = R.linear_tables(1000, 0.00085, 0.012)
        g = torch.Generator().manual_seed(1234)
        self.x0 = torch.rand(self.N, 3, self.S, self.S, generator=g) * 2 - 1
        self.cond = torch.rand(self.N, 3, self.S, self.S, generator=g) * 2 - 1
        self.audio = torch.randn(self.N, 4000, generator=g)
        self.g = g

    def step(self):
        R = self.R
        t0 = time.perf_counter()
        self.opt.zero_grad(set_to_none=True)
        noise = torch.randn(self.x0.shape, generator=self.g)
        t = torch.randint(0, 1000, (self.N,), generator=self.g)
        xt = R.add_noise(self.tab, self.x0, noise, t)
        feats = self.enc(input_values=self.audio, output_hidden_states=True).last_hidden_state      # unet_audio.py:16-18
        pred = R.unet_audio_forward(self.sd, self.cfg, xt, self.cond, feats.mean(dim=1), t)
        loss = R.mse(pred, noise)
        loss.backward()
        self.opt.step()
        return time.perf_counter() - t0


def cpu_train_baseline(cores, frames, reps=5, budget_s=100.0):
    """1 warm-up + median of `reps` cfg1 steps; clips/s in units of `frames`-frame clips."""
    b = CpuTrainCfg1(cores)
    b.step()
    times, t_end = [], time.perf_counter() + budget_s
    while len(times) < reps and (not times or time.perf_counter() < t_end):
        times.append(b.step())
    med = sorted(times)[len(times) // 2]
    return {"value": round((b.N / frames) / med, 5), "unit": "clips/s", "cores": cores, "kind": "port",
            "sample": f"cfg1 exactly: 16 images 64x64 fp32 (2 clips x 8 frames; clips/s quoted in {frames}-frame clips), "
                      f"HF wav2vec2-base + oracle torso (attention checkpointed as unet.py:309) fwd+bwd+Adam over 110.7 M "
                      f"parameters, 1 warm-up + median of {len(times)}: {med:.2f} s/step",
            "sec_per_step": round(med, 3), "reps": len(times), "cpu": _cpu_name(), "torch_threads": torch.get_num_threads()}


def cpu_sampler_baseline(cores, steps=5, budget_s=75.0):
    """test.py:57-65 on the host cores: `steps` steps of oracle model forward + sample_prev_timestep at 2 images 128x128
    (conditioning hoisted like the product's loop), extrapolated linearly to 1000 / 50 steps."""
    from oracle import restatement as R
    import lipreading_video_generation_b200 as P
    torch.set_num_threads(cores)

    class Stub(torch.nn.Module):
        def forward(self, a):
            return a["feats"]
    S, n = 128, 2
    shell = P.UNetAudio(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
                        attention_resolutions=(1, 2, 4), audio_feature_dim=768, projected_audio_dim=128, audio_encoder=Stub())
    sd = R.protocol_params(shell.state_dict(), seed=0)
    cfg = dict(image_size=S, model_channels=64, num_res_blocks=2, attention_resolutions=(1, 2, 4), channel_mult=(1, 2, 4),
               num_heads=1)
    tab = R.linear_tables(1000, 0.00085, 0.012)
    g = torch.Generator().manual_seed(1234)
    xt = torch.randn(n, 3, S, S, generator=g)
    cond = torch.rand(n, 3, S, S, generator=g) * 2 - 1
    fm = torch.randn(n, 768, generator=g)
    times = []
    t_end = time.perf_counter() + budget_s
    with torch.no_grad():
        for k in range(steps):
            if k >= 2 and time.perf_counter() > t_end:
                break
            i = 999 - k
            t0 = time.perf_counter()
            eps = R.unet_audio_forward(sd, cfg, xt, cond, fm, torch.full((n,), i))
            xt, _ = R.ddpm_step(tab, xt, eps, i, torch.randn(xt.shape, generator=g))
            times.append(time.perf_counter() - t0)
    sec = sorted(times)[len(times) // 2]
    return {"value": round(n / (1000 * sec), 6), "unit": "frames/s (1000-step DDPM)", "cores": cores, "kind": "port",
            "value_50_step_ddim": round(n / (50 * sec), 5),
            "sample": f"{len(times)} steps of oracle forward + ddpm_step at {n} images 128x128 fp32, median {sec:.2f} s/step, "
                      f"EXTRAPOLATED linearly to 1000 / 50 steps", "sec_per_step": round(sec, 3), "cpu": _cpu_name()}


# ----------------------------------------------------------------------------- GPU measurements
def _barrier(world):
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(vals, world, device):
    if world > 1:
        tt = torch.tensor(vals, device=device, dtype=torch.float64)
        torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        return tt.tolist()
    return list(vals)


def _attention_roofline(totals, steps, ms_step, sustained, which):
    dom = None
    for (name, meta), (t_ms, cnt) in totals.items():
        if not name.startswith("attention"):
            continue
        Nn, Lk, heads, ch = meta[:4]
        Lq = meta[4] if len(meta) > 4 else Lk
        flops = (4.0 if name == "attention_fwd" else 10.0) * Lq * Lk * ch * heads * Nn
        rec = dict(kernel=f"{name} L={Lk} ch={ch}", ms_per_launch=t_ms / cnt, tflops=flops / (t_ms / cnt) / 1e9,
                   ms_per_step=t_ms / steps)
        if dom is None or rec["ms_per_step"] > dom["ms_per_step"]:
            dom = rec
    if dom is None:                      # no attention launch was timed (e.g. a configuration without attention)
        return {"bound": "tensor", "achieved": None, "peak": sustained, "unit": "TFLOP/s", "frac": None, "traffic": None,
                "kernel": None, "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({which})"}
    return {"bound": "tensor", "achieved": round(dom["tflops"], 1), "peak": sustained, "unit": "TFLOP/s",
            "frac": round(dom["tflops"] / sustained, 4), "traffic": None, "kernel": dom["kernel"],
            "ms_per_launch": round(dom["ms_per_launch"], 4), "share_of_step": round(dom["ms_per_step"] / ms_step, 3),
            "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({which}); algorithmic FLOPs 4*L^2*ch*N fwd, "
                           f"10*L^2*ch*N bwd"}


def _fill_traffic(rec, n_img):
    """roofline.traffic: DRAM bytes per launch of the dominant kernel from one `ncu --set full` capture at this shape
    (profiles/r2_traffic.json), null when no capture exists for it."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            table = json.load(f)["kernels"]
        hit = table.get(f'{rec["roofline"]["kernel"]}|N={n_img}')
        if hit:
            rec["roofline"]["traffic"] = hit["traffic_bytes_per_launch"]
    except (OSError, KeyError, ValueError, TypeError):
        pass


def measure_train(steps, warmup, S, clips, frames, world, rank, local, device, verbose=False):
    """Returns the record of one train workload (all ranks run it; max over ranks)."""
    from lipreading_video_generation_b200 import _lib as L, ops
    from lipreading_video_generation_b200.parallel import broadcast_parameters
    n_img = clips * frames
    model = build_model(S, device)
    if world > 1:
        broadcast_parameters(model)
    step = TrainStep(model, S, world, device)
    x0, cond, audio = make_batch(n_img, S, 1234 + rank, device)
    hx0, hcond, haudio = make_batch(n_img, S, 1234 + rank, device, pinned=True)
    gt = torch.Generator().manual_seed(99 + rank)
    t_dev = torch.randint(0, 1000, (n_img,), generator=gt).to(device)
    t_host = torch.randint(0, 1000, (n_img,), generator=gt).pin_memory()
    for _ in range(warmup):
        step(x0, cond, audio, t_dev)
    # ---------------- timed region 1: inputs resident in HBM ("value") + roofline events on the attention kernels
    _barrier(world)
    l0 = L.lib().b2d_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # only the candidate dominant kernels carry events inside the timed region (26 launches per step); the full
    # per-class breakdown (~530 event pairs per step, which costs 2-3 ms of step time) runs in an untimed pass below
    with ClockSampler(local) as clk, ops.KernelTimer({"attention_fwd", "attention_bwd"}) as kt:
        e0.record()
        h0 = time.perf_counter()
        for _ in range(steps):
            step(x0, cond, audio, t_dev)
        host_ms = (time.perf_counter() - h0) * 1e3 / steps     # host time to enqueue one step (no sync inside)
        e1.record()
        _barrier(world)
    ms = e0.elapsed_time(e1) / steps
    launches = (L.lib().b2d_launch_count() - l0) // steps
    # ---------------- timed region 2: end to end through the public API with HOST buffers
    _barrier(world)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    # the loss of every step is copied to pinned host memory and READ on the host one step later (after the next step
    # has been enqueued), the last one before the timer stops: logging the loss does not drain the launch queue
    loss_pin = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    losses_host = []
    for i in range(steps):
        dx0 = hx0.to(device, non_blocking=True); dc = hcond.to(device, non_blocking=True)
        da = haudio.to(device, non_blocking=True); dt = t_host.to(device, non_blocking=True)
        loss = step(dx0, dc, da, dt)
        loss_pin[i % 2].copy_(loss.detach(), non_blocking=True)     # device -> host read of the step's result
        loss_ev[i % 2].record()
        if i > 0:
            loss_ev[(i - 1) % 2].synchronize()
            losses_host.append(float(loss_pin[(i - 1) % 2]))
    loss_ev[(steps - 1) % 2].synchronize()
    losses_host.append(float(loss_pin[(steps - 1) % 2]))
    assert len(losses_host) == steps
    f1.record()
    _barrier(world)
    ms_e2e = f0.elapsed_time(f1) / steps
    # ---------------- untimed diagnostic pass: per-class kernel time of one step
    with ops.KernelTimer(CLASSES) as kt_all:
        step(x0, cond, audio, t_dev)
        _barrier(world)
    ms, ms_e2e = _max_over_ranks([ms, ms_e2e], world, device)
    step.close()
    burst, sustained, hbm, which = peaks()
    by_class = {}
    for (name, meta), (t_ms, cnt) in kt_all.totals().items():
        by_class[name] = by_class.get(name, 0.0) + t_ms
        if verbose and rank == 0:
            print(f"[kernel] {name:14s} {str(meta):40s} x{cnt:3d}/step  {t_ms / cnt:8.4f} ms/launch  {t_ms:8.3f} ms/step",
                  file=sys.stderr)
    roof = _attention_roofline(kt.totals(), steps, ms, sustained, which)
    train_tflop = 3 * FWD_GFLOP_PER_IMAGE[S] * n_img / 1000.0
    h2d = sum(t.numel() * t.element_size() for t in (hx0, hcond, haudio, t_host))
    rec = {"value": round(clips * world / (ms / 1000.0), 3), "unit": "clips/s", "steps": steps, "warmup": warmup,
           "ms_per_step": round(ms, 3),
           "e2e": {"value": round(clips * world / (ms_e2e / 1000.0), 3), "unit": "clips/s", "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e, 3)},
           "gpu_launches": int(launches), "host_enqueue_ms_per_step": round(host_ms, 3), "roofline": roof,
           "step_tflops_algorithmic": round(train_tflop / (ms / 1000.0), 1),
           "step_frac_of_sustained_peak": round(train_tflop / (ms / 1000.0) / sustained, 4),
           "kernel_class_ms_per_step": {k: round(v, 3) for k, v in sorted(by_class.items())},
           "clocks": clk.summary(), "last_loss": losses_host[-1], "images_per_gpu": n_img,
           # calls the patched HF encoder handed to the stock library path over the whole run, per patch site
           # (unet_audio.stock_route_counts): {} = every patched piece ran on this library's kernels
           "w2v_stock_routes": _stock_routes()}
    del step, model, x0, cond, audio
    gc.collect()
    torch.cuda.empty_cache()
    return rec


def _stock_routes():
    import lipreading_video_generation_b200.unet_audio as UA
    return UA.stock_route_counts(reset=True)


def measure_sample(steps, warmup, S, clips, frames, world, rank, local, device):
    """sampler frames/s (BASELINE configs[3] per-GPU shape): replicas only, no collective on the data path.
    Times the 1000-step DDPM update (`value`) and, separately, a real 50-step DDIM loop."""
    from lipreading_video_generation_b200 import _lib as L, ops
    n_img = clips * frames
    model = build_model(S, device)
    x0, cond, audio = make_batch(n_img, S, 1234 + rank, device)
    hx0, hcond, haudio = make_batch(n_img, S, 1234 + rank, device, pinned=True)
    out = {}
    for mode in ("ddpm", "ddim"):
        step = SampleStep(model, S, device, mode=mode)
        step.prepare(audio)
        xt = torch.randn(n_img, 3, S, S, device=device)
        for _ in range(warmup):
            xt, _ = step(xt, cond)
        _barrier(world)
        l0 = L.lib().b2d_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(local) as clk, ops.KernelTimer({"attention_fwd", "conv2d", "groupnorm_fwd"}) as kt:
            e0.record()
            h0 = time.perf_counter()
            for _ in range(steps):
                xt, x0p = step(xt, cond)
            host_ms = (time.perf_counter() - h0) * 1e3 / steps
            e1.record()
            _barrier(world)
        ms = e0.elapsed_time(e1) / steps
        launches = (L.lib().b2d_launch_count() - l0) // steps
        kt_steps = steps
        if step.graphed is not None:
            # a replay executes the recorded forward's kernels without ABI calls: count them, and time the kernel
            # classes on two eager (un-recorded, untimed-for-the-metric) steps -- events cannot be read out of a graph
            launches += step.graphed.launches_per_replay
            diag = SampleStep(model, S, device, mode=mode, graph=False)
            diag.feats, kt_steps = step.feats, 2
            xd = xt.clone()
            xd, _ = diag(xd, cond)
            with ops.KernelTimer({"attention_fwd", "conv2d", "groupnorm_fwd"}) as kt:
                for _ in range(kt_steps):
                    xd, _ = diag(xd, cond)
            torch.cuda.synchronize()
        # end to end: conditioning + x_T from pinned host memory, wav2vec2 encoding, K steps, x0_pred back to the host
        _barrier(world)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        hxt = torch.randn(n_img, 3, S, S).pin_memory()
        hout = torch.empty(n_img, 3, S, S).pin_memory()
        f0.record()
        dc = hcond.to(device, non_blocking=True); da = haudio.to(device, non_blocking=True); dx = hxt.to(device, non_blocking=True)
        step.prepare(da)
        for _ in range(steps):
            dx, x0p = step(dx, dc)
        hout.copy_(x0p, non_blocking=True)
        f1.record()
        _barrier(world)
        ms_e2e = f0.elapsed_time(f1) / steps
        ms, ms_e2e = _max_over_ranks([ms, ms_e2e], world, device)
        out[mode] = dict(ms=ms, ms_e2e=ms_e2e, launches=launches, host_ms=host_ms, kt=kt, kt_steps=kt_steps,
                         clk=clk.summary(), graph=step.graphed is not None,
                         h2d=sum(t.numel() * t.element_size() for t in (hcond, haudio, hxt)) // steps,
                         d2h=hout.numel() * 4 // steps)
    burst, sustained, hbm, which = peaks()
    d, i = out["ddpm"], out["ddim"]
    by_class = {}
    for (name, meta), (t_ms, cnt) in d["kt"].totals().items():
        by_class[name] = by_class.get(name, 0.0) + t_ms / d["kt_steps"]
    fwd_tflop = FWD_GFLOP_PER_IMAGE[S] * n_img / 1000.0
    rec = {"value": round(n_img * world / (1000 * d["ms"] / 1000.0), 4), "unit": "frames/s (1000-step DDPM)",
           "steps": steps, "warmup": warmup, "ms_per_step": round(d["ms"], 3),
           "frames_per_s_50_step_ddim": round(n_img * world / (50 * i["ms"] / 1000.0), 3),
           "ms_per_ddim_step": round(i["ms"], 3),
           "ddim_note": "measured: a 50-step eta=0 subsequence of the same schedule run through scheduler.ddim_step",
           "e2e": {"value": round(n_img * world / (1000 * d["ms_e2e"] / 1000.0), 4), "unit": "frames/s (1000-step DDPM)",
                   "h2d_bytes_per_step": d["h2d"], "d2h_bytes_per_step": d["d2h"], "ms_per_step": round(d["ms_e2e"], 3),
                   "frames_per_s_50_step_ddim": round(n_img * world / (50 * i["ms_e2e"] / 1000.0), 3)},
           "gpu_launches": int(d["launches"]), "host_enqueue_ms_per_step": round(d["host_ms"], 3),
           "model_forward": "one CUDA graph replay per step (sampling.GraphedDenoiser)" if d["graph"] else "eager launches",
           "roofline": _attention_roofline(d["kt"].totals(), d["kt_steps"], d["ms"], sustained, which),
           "step_tflops_algorithmic": round(fwd_tflop / (d["ms"] / 1000.0), 1),
           "step_frac_of_sustained_peak": round(fwd_tflop / (d["ms"] / 1000.0) / sustained, 4),
           "kernel_class_ms_per_step": {k: round(v, 3) for k, v in sorted(by_class.items())},
           "clocks": d["clk"], "images_per_gpu": n_img}
    del model, x0, cond, audio
    gc.collect()
    torch.cuda.empty_cache()
    return rec


def measure_hbm_kernels(device):
    """Achieved HBM rate of the elementwise / reduction kernels north_star names, at the BASELINE configs[2] (q_sample,
    MSE: N = 256 images 128x128) and configs[3] (sampler update: N = 128) sizes; algorithmic bytes per SURVEY 8(d)."""
    import lipreading_video_generation_b200 as P
    burst, sustained, hbm, which = peaks()
    sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
    crit = P.MSELoss()
    out = {}

    SETS = 4          # rotating input sets: the working set of consecutive launches is 4x the 126 MB L2

    def timed(fn, reps=20):
        """ms per launch of fn(i), i = launch index: the launches are recorded into ONE CUDA graph and replayed, so the
        Python wrapper (30-40 us per call, more than these kernels run) is not in the measurement."""
        for i in range(SETS):
            fn(i)
        torch.cuda.synchronize()
        side = torch.cuda.Stream(device=device)
        graph = torch.cuda.CUDAGraph()
        keep = []         # outputs stay alive through the capture: every launch writes its own buffers (a freed output
        with torch.cuda.stream(side):          # would be handed to the next launch and stay in L2)
            with torch.cuda.graph(graph, stream=side):
                for i in range(reps):
                    keep.append(fn(i))
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        del keep, graph
        return e0.elapsed_time(e1) / reps
    with torch.no_grad():
        n = 256
        x0 = [torch.rand(n, 3, 128, 128, device=device) for _ in range(SETS)]
        eps = [torch.randn_like(x0[0]) for _ in range(SETS)]
        t = torch.randint(0, 1000, (n,), device=device)
        E = x0[0].numel()
        ms = timed(lambda i: sched.add_noise(x0[i % SETS], eps[i % SETS], t))
        out["q_sample"] = {"elements": E, "bytes": 12 * E, "ms": round(ms, 4), "gbs": round(12 * E / ms / 1e6, 1)}
        ms = timed(lambda i: crit(x0[i % SETS], eps[i % SETS]))
        out["mse_fwd"] = {"elements": E, "bytes": 8 * E, "ms": round(ms, 4), "gbs": round(8 * E / ms / 1e6, 1)}
        n = 128
        xt = [torch.randn(n, 3, 128, 128, device=device) for _ in range(SETS)]
        e = [torch.randn_like(xt[0]) for _ in range(SETS)]
        z = [torch.randn_like(xt[0]) for _ in range(SETS)]
        E = xt[0].numel()
        ms = timed(lambda i: sched.sample_prev_timestep(xt[i % SETS], e[i % SETS], 500, z=z[i % SETS]))
        out["ddpm_step"] = {"elements": E, "bytes": 20 * E, "ms": round(ms, 4), "gbs": round(20 * E / ms / 1e6, 1)}
        ms = timed(lambda i: sched.ddim_step(xt[i % SETS], e[i % SETS], 500, 480))
        out["ddim_step"] = {"elements": E, "bytes": 16 * E, "ms": round(ms, 4), "gbs": round(16 * E / ms / 1e6, 1)}
    for v in out.values():
        v["frac_of_hbm_peak"] = round(v["gbs"] / hbm, 3)
    out["peak_gbs"] = hbm
    out["note"] = ("20 launches per op replayed as one CUDA graph over 4 rotating input sets (working set 4x L2), CUDA "
                   "events around the replay; bytes are algorithmic (q_sample 12 B/elem, MSE 8, DDPM update 20, DDIM 16); "
                   "ncu per-kernel DRAM counters: profiles/r2_hbm_kernels.md")
    return out


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

    for _ in range(a.warmup):
        step(x, eps, t)
    _barrier(world)
    l0 = L.lib().b2d_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk, ops.KernelTimer({"attention_fwd", "attention_bwd"}) as kt:
        e0.record()
        for _ in range(a.steps):
            step(x, eps, t)
        e1.record()
        _barrier(world)
    ms = e0.elapsed_time(e1) / a.steps
    launches = (L.lib().b2d_launch_count() - l0) // a.steps
    _barrier(world)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(a.steps):
        loss = step(hx.to(device, non_blocking=True), heps.to(device, non_blocking=True), ht.to(device, non_blocking=True))
        loss_host = loss.item()
    f1.record()
    _barrier(world)
    ms_e2e = f0.elapsed_time(f1) / a.steps
    with ops.KernelTimer(CLASSES) as kt_all:
        step(x, eps, t)
        _barrier(world)
    ms, ms_e2e = _max_over_ranks([ms, ms_e2e], world, device)
    if world > 1:
        torch.distributed.destroy_process_group()
    if rank != 0:
        return
    burst, sustained, hbm, which = peaks()
    by_class = {}
    for (name, meta), (t_ms, cnt) in kt_all.totals().items():
        by_class[name] = by_class.get(name, 0.0) + t_ms
    h2d = sum(v.numel() * v.element_size() for v in (hx, heps, ht))
    out = {"metric": "denoiser_train_clips_per_s", "value": round(clips / (ms / 1000.0), 4), "unit": "clips/s",
           "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
           "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": config,
           "seq_parallel_bytes_per_step": (sp.bytes_moved // (a.warmup + 2 * a.steps + 1)) if sp is not None else 0,
           "e2e": {"value": round(clips / (ms_e2e / 1000.0), 4), "unit": "clips/s", "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e, 3)},
           "gpu_launches": int(launches), "roofline": _attention_roofline(kt.totals(), a.steps, ms, sustained, which),
           "kernel_class_ms_per_step": {k: round(v, 3) for k, v in sorted(by_class.items())},
           "cpu_baseline": None, "clocks": clk.summary(), "last_loss": loss_host}
    print(json.dumps(out))


def _config(name, clips, frames, S, kind, world):
    cfg = {"workload": f"{name}: denoiser {kind} step, {clips} clips x {frames} frames {S}x{S} per GPU "
                       f"(N={clips * frames} images folded into the batch), UNetAudio train.py config + wav2vec2-base",
           "clips_per_gpu": clips, "frames": frames, "image": S, "parallelism": f"dp{world}",
           "l2_policy": "inputs+activations >> L2 (126 MB): >10 GB touched per step"}
    if kind == "train3d":
        cfg["workload"] = (f"{name}: UNetModel(dims=3) train step on {clips} clip(s) (195, {frames}, {S}, {S}) per GPU, "
                           f"train.py torso config (64 ch, mult (1,2,4), 2 res blocks, attention at all levels over "
                           f"L = T*S*S = {frames * S * S} positions), MSE + backward + Adam, no audio encoder")
    return cfg


def main_reference(a, frames, config, cores, kind):
    """--impl reference: the reference's CPU path on the host cores (oracle port; /root/reference does not travel to
    the GPU box).  Train workloads: each step is ONE cfg1 train step (16 images 64x64, fwd+bwd+Adam incl. wav2vec2) --
    a bounded sample of the workload; W warm-up + K timed steps, capped by a wall-clock budget with the real number
    of timed steps reported.  Sample workloads: steps of oracle forward + ddpm_step at 2 images 128x128."""
    budget = float(os.environ.get("B2D_REF_BUDGET_S", "240"))
    t_start = time.perf_counter()
    if kind == "sample":
        r = cpu_sampler_baseline(cores, steps=max(1, min(a.steps, 5)))
        line = {"impl": "reference", "metric": "sampler_frames_per_s", "value": r["value"], "unit": r["unit"],
                "n_gpus": a.gpus, "steps": max(1, min(a.steps, 5)), "steps_requested": a.steps, "warmup": 0,
                "ms_per_step": r["sec_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config, "cpu_baseline": r,
                "e2e": {"value": r["value"], "unit": r["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return
    b = CpuTrainCfg1(cores)
    warm = 0
    for _ in range(max(1, a.warmup)):
        b.step()
        warm += 1
        if time.perf_counter() - t_start > 0.25 * budget:
            break
    times = []
    for _ in range(a.steps):
        times.append(b.step())
        if time.perf_counter() - t_start > budget:
            break
    mean = sum(times) / len(times)
    v = (b.N / frames) / mean
    cb = {"value": v, "unit": "clips/s", "cores": cores, "kind": "port", "cpu": _cpu_name(),
          "sample": f"each step = cfg1 exactly: 16 images 64x64 fp32 (clips/s quoted in {frames}-frame clips), HF wav2vec2-base + "
                    f"oracle torso (attention checkpointed as unet.py:309) fwd+bwd+Adam over 110.7 M parameters; "
                    f"{warm} warm-up + {len(times)} timed steps, mean {mean:.2f} s/step"}
    print(json.dumps({"impl": "reference", "metric": "denoiser_train_clips_per_s", "value": v, "unit": "clips/s",
                      "n_gpus": a.gpus, "steps": len(times), "steps_requested": a.steps, "warmup": warm,
                      "ms_per_step": mean * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                      "dtype": "f32", "data": "synthetic", "config": config, "cpu_baseline": cb,
                      "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="train64", choices=list(WORKLOADS))
    ap.add_argument("--no-extras", action="store_true", help="train64 only: skip the sample128 / train128 / HBM-kernel extras")
    a = ap.parse_args()
    clips, frames, S, kind = WORKLOADS[a.workload]
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = os.cpu_count() or 1
    config = _config(a.workload, clips, frames, S, kind, world)
    no_cpu = os.environ.get("B2D_BENCH_NO_CPU") == "1"      # profiling / development runs: skip the host-side legs

    if a.impl == "reference":
        if rank == 0:
            main_reference(a, frames, config, cores, kind)
        return

    assert torch.cuda.is_available(), "bench.py needs a B200; there is no CPU fallback for the product path"
    torch.backends.cudnn.benchmark = True      # third-party wav2vec2 (HF / cuDNN) only: let cuDNN pick its conv kernels
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    from lipreading_video_generation_b200 import _lib as L
    L.check(L.lib().b2d_check_device(), "b2d_check_device")
    if kind == "train3d":
        return main_train3d(a, S, clips, frames, world, rank, local, device, config, cores)

    if kind == "sample":
        rec = measure_sample(a.steps, a.warmup, S, clips, frames, world, rank, local, device)
        _fill_traffic(rec, clips * frames)
        out = {"metric": "sampler_frames_per_s", "n_gpus": world, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": config, **rec}
        if rank == 0:
            out["cpu_baseline"] = None if (no_cpu or world > 1) else cpu_sampler_baseline(cores)
            print(json.dumps(out))
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    rec = measure_train(a.steps, a.warmup, S, clips, frames, world, rank, local, device, verbose=True)
    _fill_traffic(rec, clips * frames)
    out = {"metric": "denoiser_train_clips_per_s", "n_gpus": world, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": config, **rec}
    extras = {}
    if a.workload == "train64" and not a.no_extras:
        ks = max(3, min(a.steps, 5))
        c2, f2, s2, _ = WORKLOADS["sample128"]
        r = measure_sample(ks, 3, s2, c2, f2, world, rank, local, device)
        r["config"] = _config("sample128", c2, f2, s2, "sample", world)
        _fill_traffic(r, c2 * f2)
        extras["sample128"] = r
        c3, f3, s3, _ = WORKLOADS["train128"]
        r = measure_train(max(3, min(a.steps, 3)), 3, s3, c3, f3, world, rank, local, device)
        r["config"] = _config("train128", c3, f3, s3, "train", world)
        _fill_traffic(r, c3 * f3)
        extras["train128"] = r
        if rank == 0:
            extras["hbm_kernels"] = measure_hbm_kernels(device)
    if world > 1:
        torch.distributed.barrier()
    if rank == 0:
        if no_cpu or world > 1:
            # the CPU baseline is timed at N = 1 only (one process owns the host cores; under torchrun the other ranks
            # spin in the barrier and OMP_NUM_THREADS is 1)
            out["cpu_baseline"] = {"value": None, "unit": "clips/s", "cores": cores, "kind": "port",
                                   "sample": "skipped (B2D_BENCH_NO_CPU=1)" if no_cpu else "timed at N = 1 only"}
        else:
            out["cpu_baseline"] = cpu_train_baseline(cores, frames)
            if "sample128" in extras:
                extras["sample128"]["cpu_baseline"] = cpu_sampler_baseline(cores)
        if extras:
            out["extra"] = extras
        print(json.dumps(out))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
