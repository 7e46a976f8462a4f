"""Sampling loop -- the recipe of video-generation/diffusion/test.py:51-83 (`sample_images`) without its
per-step allocator flush, autograd graph and recomputation of the loop-invariant conditioning."""
from __future__ import annotations

import torch


@torch.no_grad()
def sample_images(model, scheduler, img_cond, audio_cond, n_timesteps, xt=None, noises=None, callback=None,
                  ddim_steps=None, graph=False):
    """Returns (x_final, x0_pred).  `xt` / `noises[i]` inject the initial noise and the per-step z so that a
    trajectory can be compared step by step with the reference; `callback(i, xt, x0_pred)` replaces the
    PNG dumps of test.py:67-81.  With ddim_steps=K a strided eta=0 DDIM trajectory is run instead.  graph=True records
    the model forward once as a CUDA graph and replays it every step (GraphedDenoiser; UNetAudio only)."""
    model.eval()
    device = img_cond.device
    N = img_cond.shape[0]
    if xt is None:
        xt = torch.randn((N, model.out_channels, model.image_size, model.image_size), device=device)
    feats = model.encode_audio(audio_cond) if hasattr(model, "encode_audio") else None   # hoisted: loop invariant
    x0_pred = None
    if graph:
        if feats is None:
            raise RuntimeError("graph=True needs the audio-conditioned model (UNetAudio)")
        graphed = GraphedDenoiser(model, img_cond, feats)

        def denoise(xt_, i_, t_):
            return graphed(xt_, i_)
    else:
        def denoise(xt_, i_, t_):
            return model(xt_, img_cond, audio_cond, t_, audio_features=feats) if feats is not None else model(xt_, t_)
    if ddim_steps is None:
        steps = list(reversed(range(n_timesteps)))
        for k, i in enumerate(steps):
            t = torch.full((N,), i, device=device, dtype=torch.long)
            noise_pred = denoise(xt, i, t)
            z = noises[k] if noises is not None else None
            xt, x0_pred = scheduler.sample_prev_timestep(xt, noise_pred, i, z=z)
            if callback is not None:
                callback(i, xt, x0_pred)
    else:
        seq = [int(round(v)) for v in torch.linspace(n_timesteps - 1, 0, ddim_steps).tolist()]
        for k, i in enumerate(seq):
            t = torch.full((N,), i, device=device, dtype=torch.long)
            noise_pred = denoise(xt, i, t)
            prev = seq[k + 1] if k + 1 < len(seq) else -1
            xt, x0_pred = scheduler.ddim_step(xt, noise_pred, i, prev)
            if callback is not None:
                callback(i, xt, x0_pred)
    return xt, x0_pred


class GraphedDenoiser:
    """The sampler's model forward as ONE CUDA graph launch (static workspace + graph capture).

    test.py:57-65 calls the model 1000 times on tensors of one shape with loop-invariant conditioning, so the ~330
    kernel launches of a forward (at batch 1 each far shorter than its launch cost) are recorded once and replayed.
    The graph reads its inputs from static buffers: `x` (noisy frames, copied in per step), `t` (timestep, filled per
    step), and private copies of the reference image and the time-averaged audio features; every intermediate and the
    output live in the graph's private memory pool, and the per-pass zero-filled accumulators come from an arena owned
    by this object (cleared by a memset node inside the graph).  The returned tensor is the graph's own output buffer:
    it is overwritten by the next call, so consume it first (the sampler update does).

    The recorded kernels read cached operand copies of the weights: after changing parameters call `update()`, which
    records the forward again."""

    def __init__(self, model, img_cond, feats, warmup=2):
        from .engine import _ZeroArena
        if not img_cond.is_cuda:
            raise RuntimeError("GraphedDenoiser needs CUDA tensors on an sm_100a device (no CPU fallback)")
        self.model = model.eval()
        self.eng = model.engine()
        dev = img_cond.device
        N = img_cond.shape[0]
        self.x = torch.zeros((N, model.out_channels, model.image_size, model.image_size), device=dev)
        self.t = torch.zeros((N,), device=dev, dtype=torch.long)
        self.cond = img_cond.detach().clone()
        self.feats = feats.detach().float().clone()
        self._arena = _ZeroArena()
        self.graph = self.out = None
        self.launches_per_replay = 0
        self._warmup = max(2, warmup)
        self._capture()

    def _capture(self):
        from . import _lib as L
        dev = self.x.device
        self.graph = self.out = None                  # drop the previous recording (and its memory pool) first
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(self._warmup):             # pass 1 sizes the arena and packs the weights, pass 2 runs as recorded
                self._run()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        l0 = L.lib().b2d_launch_count()
        with torch.no_grad(), torch.cuda.graph(graph):
            out = self._run()
        self.launches_per_replay = L.lib().b2d_launch_count() - l0
        self.graph, self.out = graph, out

    def _run(self):
        from . import ops
        eng = self.eng
        keep, eng._fwd_arena = eng._fwd_arena, self._arena
        timer, ops.KernelTimer.active = ops.KernelTimer.active, None       # timing events cannot be read back from a graph
        try:
            out, _ = eng.forward(self.x, self.cond, self.feats, self.t, train=False)
        finally:
            eng._fwd_arena = keep
            ops.KernelTimer.active = timer
        return out

    def update(self):
        """Parameters changed: the recorded kernels read cached operand copies (packed fp16 weights, the concatenated
        embedding projections), so the forward is recorded again from the current parameters."""
        self.eng.invalidate()
        self._capture()

    def set_conditioning(self, img_cond, feats):
        """Other conditioning of the same shape (next clip): copied into the buffers the recording reads."""
        self.cond.copy_(img_cond)
        self.feats.copy_(feats)

    @torch.no_grad()
    def __call__(self, xt, i):
        self.x.copy_(xt)
        self.t.fill_(int(i))
        self.graph.replay()
        return self.out
