"""Sampling loop -- the recipe of video-generation/diffusion/test.py:51-83 (`sample_images`) without its
per-step allocator flush, autograd graph and recomputation of the loop-invariant conditioning."""
from __future__ import annotations

import torch


@torch.no_grad()
def sample_images(model, scheduler, img_cond, audio_cond, n_timesteps, xt=None, noises=None, callback=None,
                  ddim_steps=None):
    """Returns (x_final, x0_pred).  `xt` / `noises[i]` inject the initial noise and the per-step z so that a
    trajectory can be compared step by step with the reference; `callback(i, xt, x0_pred)` replaces the
    PNG dumps of test.py:67-81.  With ddim_steps=K a strided eta=0 DDIM trajectory is run instead."""
    model.eval()
    device = img_cond.device
    N = img_cond.shape[0]
    if xt is None:
        xt = torch.randn((N, model.out_channels, model.image_size, model.image_size), device=device)
    feats = model.encode_audio(audio_cond) if hasattr(model, "encode_audio") else None   # hoisted: loop invariant
    x0_pred = None
    if ddim_steps is None:
        steps = list(reversed(range(n_timesteps)))
        for k, i in enumerate(steps):
            t = torch.full((N,), i, device=device, dtype=torch.long)
            noise_pred = model(xt, img_cond, audio_cond, t, audio_features=feats) if feats is not None else model(xt, t)
            z = noises[k] if noises is not None else None
            xt, x0_pred = scheduler.sample_prev_timestep(xt, noise_pred, i, z=z)
            if callback is not None:
                callback(i, xt, x0_pred)
    else:
        seq = [int(round(v)) for v in torch.linspace(n_timesteps - 1, 0, ddim_steps).tolist()]
        for k, i in enumerate(seq):
            t = torch.full((N,), i, device=device, dtype=torch.long)
            noise_pred = model(xt, img_cond, audio_cond, t, audio_features=feats) if feats is not None else model(xt, t)
            prev = seq[k + 1] if k + 1 < len(seq) else -1
            xt, x0_pred = scheduler.ddim_step(xt, noise_pred, i, prev)
            if callback is not None:
                callback(i, xt, x0_pred)
    return xt, x0_pred
