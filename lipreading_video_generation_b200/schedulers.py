"""Noise schedulers -- drop-ins for video-generation/diffusion/linear_noise_scheduler.py and
noise_scheduler.py.  Same constructors, public table attributes, `add_noise` and
`sample_prev_timestep` signatures and return values.

Tables are built with the reference's own torch ops on the host (bit-exact fp32), uploaded ONCE per
device (the reference re-uploads them on every call, linear_noise_scheduler.py:35-36,57-68), and the
arithmetic is one fused sm_100a kernel per call.  Per-step scalar coefficients of the samplers are
gathered on the host from the CPU tables, which also removes the reference's `if t == 0` device sync.

Extras that the reference does not have (documented, opt-in): `z=` to inject the sampler noise for
trajectory parity tests, and `ddim_step` (eta = 0; parity unpinned).
"""
from __future__ import annotations

import math

import torch

from . import _lib as L


def _t_index(t) -> int:
    if isinstance(t, torch.Tensor):
        if t.numel() != 1:
            raise RuntimeError("sample_prev_timestep expects a single timestep (the reference's tables are indexed "
                               "with a 1-element tensor and broadcast over the batch)")
        return int(t.reshape(()).item())
    return int(t)


class _DeviceTables:
    def __init__(self):
        self._dev = {}

    def get(self, device, **tables):
        key = str(device)
        if key not in self._dev:
            self._dev[key] = {k: v.to(device).contiguous() for k, v in tables.items()}
        return self._dev[key]


def _launch_step(name, xt, noise_pred, z, need_z, *scalars_and_flags):
    L.require_cuda(xt, noise_pred)
    xt_c, e_c = xt.contiguous().float(), noise_pred.contiguous().float()
    if need_z and z is None:
        z = torch.randn_like(xt_c)
    z_c = z.contiguous().float() if z is not None else None
    a = torch.empty_like(xt_c)
    b = torch.empty_like(xt_c)
    fn = getattr(L.lib(), name)
    L.check(fn(xt_c.data_ptr(), e_c.data_ptr(), L.ptr(z_c), *scalars_and_flags, a.data_ptr(), b.data_ptr(),
               xt_c.numel(), L.stream_ptr()), name)
    return a, b


class LinearNoiseScheduler:
    r"""linear_noise_scheduler.py:6-76"""

    def __init__(self, num_timesteps, beta_start, beta_end):
        self.num_timesteps = num_timesteps
        self.beta_start = beta_start
        self.beta_end = beta_end
        self.betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_timesteps) ** 2
        self.alphas = 1. - self.betas
        self.alpha_cum_prod = torch.cumprod(self.alphas, dim=0)
        self.sqrt_alpha_cum_prod = torch.sqrt(self.alpha_cum_prod)
        self.sqrt_one_minus_alpha_cum_prod = torch.sqrt(1 - self.alpha_cum_prod)
        self._tabs = _DeviceTables()

    # -- q_sample ---------------------------------------------------------
    def add_noise(self, original, noise, t):
        """linear_noise_scheduler.py:24-46.  original/noise (B, ...) fp32 CUDA, t (B,) integer tensor."""
        L.require_cuda(original, noise)
        B = original.shape[0]
        tabs = self._tabs.get(original.device, a=self.sqrt_alpha_cum_prod, b=self.sqrt_one_minus_alpha_cum_prod)
        x0 = original.contiguous().float()
        eps = noise.contiguous().float()
        tt = t.to(device=original.device, dtype=torch.int64).reshape(B).contiguous()
        out = torch.empty_like(x0)
        per = x0.numel() // B if B else 0
        flag = torch.zeros(1, device=original.device, dtype=torch.int32)
        L.check(L.lib().b2d_q_sample(x0.data_ptr(), eps.data_ptr(), tt.data_ptr(), tabs["a"].data_ptr(),
                                     tabs["b"].data_ptr(), self.num_timesteps, out.data_ptr(), B, per,
                                     flag.data_ptr(), L.stream_ptr()), "b2d_q_sample")
        self._last_index_flag = flag   # checked lazily: no device sync on the hot path
        return out

    def check_indices(self):
        """Raise IndexError (as the reference does eagerly) if the last add_noise saw t outside [0, T)."""
        f = getattr(self, "_last_index_flag", None)
        if f is not None and int(f.item()) != 0:
            raise IndexError(f"timestep index out of bounds for schedule of size {self.num_timesteps}")

    # -- DDPM posterior step ------------------------------------------------
    def _coef(self, t):
        i = _t_index(t)
        if not (-self.num_timesteps <= i < self.num_timesteps):
            raise IndexError(f"index {i} is out of bounds for dimension 0 with size {self.num_timesteps}")
        return i

    def sample_prev_timestep(self, xt, noise_pred, t, z=None):
        """linear_noise_scheduler.py:48-76.  Returns (x_{t-1}, x0_pred)."""
        i = self._coef(t)
        s1m = float(self.sqrt_one_minus_alpha_cum_prod[i])
        sqrt_acp = float(torch.sqrt(self.alpha_cum_prod[i]))
        beta = float(self.betas[i])
        sqrt_alpha = float(torch.sqrt(self.alphas[i]))
        if i == 0:
            sigma, use_noise = 0.0, 0
        else:
            variance = (1 - self.alpha_cum_prod[i - 1]) / (1.0 - self.alpha_cum_prod[i])
            variance = variance * self.betas[i]
            sigma, use_noise = float(variance ** 0.5), 1
        return _launch_step("b2d_ddpm_step", xt, noise_pred, z, bool(use_noise), s1m, sqrt_acp, beta, sqrt_alpha,
                            sigma, use_noise)

    def ddim_step(self, xt, noise_pred, t, t_prev):
        """DDIM eta=0 (Song et al. eq. 12) on this schedule's alpha_cum_prod; t_prev < 0 means the final step.
        Not in the reference: parity unpinned (checked against the CPU restatement in oracle/)."""
        i = self._coef(t)
        acp_prev = self.alpha_cum_prod[t_prev] if t_prev >= 0 else torch.tensor(1.0)
        return _ddim(self, xt, noise_pred, float(self.sqrt_one_minus_alpha_cum_prod[i]),
                     float(self.sqrt_alpha_cum_prod[i]), float(torch.sqrt(acp_prev)), float(torch.sqrt(1 - acp_prev)))


def _ddim(self, xt, noise_pred, s1m_t, sqrt_acp_t, sqrt_acp_prev, s1m_prev):
    L.require_cuda(xt, noise_pred)
    xt_c, e_c = xt.contiguous().float(), noise_pred.contiguous().float()
    a, b = torch.empty_like(xt_c), torch.empty_like(xt_c)
    L.check(L.lib().b2d_ddim_step(xt_c.data_ptr(), e_c.data_ptr(), s1m_t, sqrt_acp_t, sqrt_acp_prev, s1m_prev,
                                  a.data_ptr(), b.data_ptr(), xt_c.numel(), L.stream_ptr()), "b2d_ddim_step")
    return a, b


class LinearNoiseSchedulerV2:
    """linear_noise_scheduler.py:79-101 (the sampler test.py:111 uses)."""

    def __init__(self, num_timesteps, beta_start=0.0001, beta_end=0.01):
        self.num_timesteps = num_timesteps
        self.beta_start = beta_start
        self.beta_end = beta_end
        self.betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_timesteps) ** 2
        self.alphas = 1. - self.betas
        self.alpha_cum_prod = torch.cumprod(self.alphas, dim=0)
        self.sqrt_alpha_cum_prod = torch.sqrt(self.alpha_cum_prod)
        self.sqrt_one_minus_alpha_cum_prod = torch.sqrt(1 - self.alpha_cum_prod)

    def sample_prev_timestep(self, xt, noise_pred, t, z=None):
        i = _t_index(t)
        if not (-self.num_timesteps <= i < self.num_timesteps):
            raise IndexError(f"index {i} is out of bounds for dimension 0 with size {self.num_timesteps}")
        s1m = float(self.sqrt_one_minus_alpha_cum_prod[i])
        sqrt_alpha = float(torch.sqrt(self.alphas[i]))
        sqrt_acp = float(self.sqrt_alpha_cum_prod[i])
        sigma = float(torch.sqrt((1 - self.alpha_cum_prod[i]) * self.betas[i]))
        return _launch_step("b2d_ddpm_v2_step", xt, noise_pred, z, True, s1m, sqrt_alpha, sqrt_acp, sigma)

    def ddim_step(self, xt, noise_pred, t, t_prev):
        i = _t_index(t)
        acp_prev = self.alpha_cum_prod[t_prev] if t_prev >= 0 else torch.tensor(1.0)
        return _ddim(self, xt, noise_pred, float(self.sqrt_one_minus_alpha_cum_prod[i]),
                     float(self.sqrt_alpha_cum_prod[i]), float(torch.sqrt(acp_prev)), float(torch.sqrt(1 - acp_prev)))


class CosineNoiseScheduler:
    """noise_scheduler.py:4-29"""

    def __init__(self, num_timesteps, s=0.008):
        self.num_timesteps = num_timesteps
        self.s = s
        self.timesteps = torch.arange(num_timesteps, dtype=torch.float32) / num_timesteps
        self.alphas_cumprod = torch.cos(((self.timesteps + self.s) / (1 + self.s)) * math.pi * 0.5) ** 2
        self.sqrt_alphas_cumprod = torch.sqrt(self.alphas_cumprod)
        self.sqrt_one_minus_alphas_cumprod = torch.sqrt(1 - self.alphas_cumprod)

    def sample_prev_timestep(self, xt, noise_pred, t, z=None):
        i = _t_index(t)
        if not (-self.num_timesteps <= i < self.num_timesteps):
            raise IndexError(f"index {i} is out of bounds for dimension 0 with size {self.num_timesteps}")
        sqrt_alpha = float(self.sqrt_alphas_cumprod[i])
        s1m = float(self.sqrt_one_minus_alphas_cumprod[i])
        variance = 1e-5
        if i > 0:
            variance = self.alphas_cumprod[i - 1] * (1 - self.alphas_cumprod[i]) / (1 - self.alphas_cumprod[i - 1])
        sigma = float(torch.sqrt(torch.as_tensor(variance, dtype=torch.float32)))
        use_noise = 1 if i > 0 else 0
        # the reference draws z even at t == 0 (and discards it); keep RNG consumption identical
        if z is None:
            z = torch.randn_like(xt)
        return _launch_step("b2d_cosine_step", xt, noise_pred, z, bool(use_noise), s1m, sqrt_alpha, sigma, use_noise)
