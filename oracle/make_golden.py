"""Generate tests/golden/*.pt from the REAL reference (imported from /root/reference) and pin
oracle/restatement.py against it.  Run in the build container:  python oracle/make_golden.py

Fixtures hold inputs + the reference's outputs only (weights are regenerated from a seed with
restatement.fill_params, applied identically to the reference model here and to the product in tests).
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_import, restatement as R  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

TINY = dict(image_size=32, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=1,
            attention_resolutions=(1, 2), channel_mult=(1, 2), dropout=0.0, audio_feature_dim=768,
            projected_audio_dim=128)


def tiny_cfg():
    return dict(image_size=TINY["image_size"], model_channels=TINY["model_channels"],
                num_res_blocks=TINY["num_res_blocks"], attention_resolutions=TINY["attention_resolutions"],
                channel_mult=TINY["channel_mult"], num_heads=1)


def make_inputs(N, S, seed, T):
    g = torch.Generator().manual_seed(seed)
    x0 = torch.rand(N, 3, S, S, generator=g) * 2 - 1
    cond = torch.rand(N, 3, S, S, generator=g) * 2 - 1
    feats = torch.randn(N, 12, 768, generator=g) * 0.5       # stands for wav2vec2 last_hidden_state
    eps = torch.randn(N, 3, S, S, generator=g)
    t = torch.randint(0, T, (N,), generator=g)
    return x0, cond, feats, eps, t


class _StubEncoder(torch.nn.Module):
    """Replaces Wav2Vec2Encoder in the reference model: returns the provided (N,12,768) features."""

    def forward(self, audio):
        return audio["feats"]


def schedules(ref):
    out = {}
    for name, cls, args in [("lin100", ref.linear_noise_scheduler.LinearNoiseScheduler, (100, 0.00085, 0.012)),
                            ("lin1000", ref.linear_noise_scheduler.LinearNoiseScheduler, (1000, 0.00085, 0.012)),
                            ("v2_500", ref.linear_noise_scheduler.LinearNoiseSchedulerV2, (500, 5e-5, 0.015))]:
        s = cls(*args)
        tabs = dict(betas=s.betas, alphas=s.alphas, alpha_cum_prod=s.alpha_cum_prod,
                    sqrt_alpha_cum_prod=s.sqrt_alpha_cum_prod,
                    sqrt_one_minus_alpha_cum_prod=s.sqrt_one_minus_alpha_cum_prod)
        mine = R.linear_tables(*args)
        for k in tabs:
            assert torch.equal(tabs[k], mine[k]), (name, k)
        out[name] = dict(args=args, **{k: v.clone() for k, v in tabs.items()})
    c = ref.noise_scheduler.CosineNoiseScheduler(1000)
    mine = R.cosine_tables(1000)
    assert torch.equal(c.alphas_cumprod, mine["alphas_cumprod"])
    out["cos1000"] = dict(args=(1000,), alphas_cumprod=c.alphas_cumprod.clone(),
                          sqrt_alphas_cumprod=c.sqrt_alphas_cumprod.clone(),
                          sqrt_one_minus_alphas_cumprod=c.sqrt_one_minus_alphas_cumprod.clone())
    return out


def sampler_vectors(ref):
    """Reference sampler outputs with the RNG noise captured (torch.manual_seed before each call)."""
    g = torch.Generator().manual_seed(11)
    xt = torch.randn(2, 3, 16, 16, generator=g)
    e = torch.randn(2, 3, 16, 16, generator=g)
    out = dict(xt=xt, eps=e, cases=[])
    lin = ref.linear_noise_scheduler.LinearNoiseScheduler(1000, 0.00085, 0.012)
    v2 = ref.linear_noise_scheduler.LinearNoiseSchedulerV2(500, 5e-5, 0.015)
    cos = ref.noise_scheduler.CosineNoiseScheduler(1000)
    for kind, sch, ts in [("ddpm", lin, [0, 1, 500, 999]), ("v2", v2, [0, 250, 499]), ("cos", cos, [0, 1, 500, 998])]:
        for t in ts:
            tt = torch.tensor([t])
            torch.manual_seed(100 + t)
            a, b = sch.sample_prev_timestep(xt, e, tt)
            torch.manual_seed(100 + t)
            z = torch.randn(xt.shape) if kind == "ddpm" else torch.randn_like(xt)
            # pin the restatement (bit-exact)
            if kind == "ddpm":
                ra, rb = R.ddpm_step(R.linear_tables(1000, 0.00085, 0.012), xt, e, t, z)
            elif kind == "v2":
                ra, rb = R.ddpm_v2_step(R.linear_tables(500, 5e-5, 0.015), xt, e, t, z)
            else:
                ra, rb = R.cosine_step(R.cosine_tables(1000), xt, e, t, z)
            assert torch.equal(a, ra) and torch.equal(b, rb), (kind, t)
            out["cases"].append(dict(kind=kind, t=t, z=z, x_prev=a.clone(), second=b.clone()))
    return out


def q_sample_vectors(ref):
    g = torch.Generator().manual_seed(12)
    x0 = torch.rand(5, 3, 16, 16, generator=g) * 2 - 1
    e = torch.randn(5, 3, 16, 16, generator=g)
    t = torch.tensor([0, 1, 500, 998, 999])
    lin = ref.linear_noise_scheduler.LinearNoiseScheduler(1000, 0.00085, 0.012)
    xt = lin.add_noise(x0, e, t)
    assert torch.equal(xt, R.add_noise(R.linear_tables(1000, 0.00085, 0.012), x0, e, t))
    return dict(x0=x0, eps=e, t=t, xt=xt)


def unet_vectors(ref):
    """Tiny UNetAudio: forward, MSE loss and parameter gradients from the real reference classes."""
    torch.manual_seed(0)
    model = ref.unet_audio.UNetAudio(**TINY)
    model.audio_encoder = _StubEncoder()
    sd = {k: v for k, v in model.state_dict().items()}
    new = R.fill_params(sd, seed=1)
    model.load_state_dict(new)
    model.train()                                   # dropout = 0.0, so train mode is deterministic
    N, S, T = 2, TINY["image_size"], 1000
    x0, cond, feats, eps, t = make_inputs(N, S, 21, T)
    lin = ref.linear_noise_scheduler.LinearNoiseScheduler(T, 0.00085, 0.012)
    xt = lin.add_noise(x0, eps, t)
    pred = model(xt, cond, {"feats": feats}, t)
    loss = torch.nn.MSELoss()(pred, eps)
    loss.backward()
    grads = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
    # ---- pin the restatement: forward, loss, every gradient
    sdr = {k: v.clone().requires_grad_(True) for k, v in new.items()}
    rp = R.unet_audio_forward(sdr, tiny_cfg(), xt, cond, feats.mean(dim=1), t)
    rl = R.mse(rp, eps)
    rl.backward()
    assert (rp - pred).abs().max().item() <= 2e-5, (rp - pred).abs().max().item()
    assert abs(rl.item() - loss.item()) <= 1e-6
    worst = 0.0
    for k, gv in grads.items():
        rg = sdr[k].grad
        rel = ((rg - gv).norm() / (gv.norm() + 1e-12)).item()
        worst = max(worst, rel)
    assert worst <= 1e-4, worst
    print(f"restatement pinned: max|pred diff| {(rp - pred).abs().max().item():.2e}, worst grad rel {worst:.2e}")
    keep = ["out.2.weight", "input_blocks.0.0.weight", "cond_conv_in.weight", "audio_transformer.transform.0.weight",
            "time_embed.0.weight", "middle_block.1.qkv.weight", "output_blocks.1.2.conv.weight",
            "input_blocks.2.0.op.weight", "input_blocks.1.0.emb_layers.1.weight", "output_blocks.0.0.skip_connection.weight"]
    return dict(cfg=dict(TINY), seed=1, x0=x0, cond=cond, feats=feats, eps=eps, t=t, xt=xt, pred=pred.detach(),
                loss=loss.detach(), grad_norms={k: v.norm().item() for k, v in grads.items()},
                grads={k: grads[k] for k in keep})


TINY3D = dict(image_size=16, in_channels=64, model_channels=64, out_channels=3, num_res_blocks=1,
              attention_resolutions=(1, 2), dropout=0.0, channel_mult=(1, 2), dims=3, num_heads=1)


def unet3d_vectors(ref):
    """Tiny UNetModel(dims=3): forward, MSE loss and parameter gradients from the real
    reference class on a (2, 64, 4, 16, 16) input; pins the restatement's conv_nd / temporal paths."""
    torch.manual_seed(0)
    model = ref.unet.UNetModel(**TINY3D)
    new = R.fill_params({k: v for k, v in model.state_dict().items()}, seed=2)
    model.load_state_dict(new)
    model.train()
    g = torch.Generator().manual_seed(31)
    B, C, T, S = 2, TINY3D["in_channels"], 4, TINY3D["image_size"]
    x = torch.randn(B, C, T, S, S, generator=g)
    target = torch.randn(B, 3, T, S, S, generator=g)
    t = torch.tensor([17, 903])
    pred = model(x, t)
    loss = torch.nn.MSELoss()(pred, target)
    loss.backward()
    grads = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
    cfg = dict(model_channels=64, num_res_blocks=1, attention_resolutions=(1, 2), channel_mult=(1, 2), num_heads=1)
    sdr = {k: v.clone().requires_grad_(True) for k, v in new.items()}
    rp = R.unet_forward(sdr, cfg, x, t)
    rl = R.mse(rp, target)
    rl.backward()
    assert (rp - pred).abs().max().item() <= 2e-5, (rp - pred).abs().max().item()
    worst = max(((sdr[k].grad - gv).norm() / (gv.norm() + 1e-12)).item() for k, gv in grads.items())
    assert worst <= 1e-4, worst
    print(f"dims=3 restatement pinned: max|pred diff| {(rp - pred).abs().max().item():.2e}, worst grad rel {worst:.2e}")
    keep = ["out.2.weight", "input_blocks.0.0.weight", "input_blocks.2.0.op.weight", "output_blocks.1.2.conv.weight",
            "middle_block.1.qkv.weight", "input_blocks.1.0.in_layers.2.weight", "output_blocks.0.0.skip_connection.weight",
            "input_blocks.1.0.emb_layers.1.weight", "time_embed.0.weight", "out.2.bias"]
    return dict(cfg=dict(TINY3D), oracle_cfg=cfg, seed=2, x=x, target=target, t=t, pred=pred.detach(), loss=loss.detach(),
                grad_norms={k: v.norm().item() for k, v in grads.items()}, grads={k: grads[k] for k in keep})


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    ref = ref_import.load()
    os.makedirs(GOLD, exist_ok=True)
    torch.save(schedules(ref), os.path.join(GOLD, "schedules.pt"))
    torch.save(sampler_vectors(ref), os.path.join(GOLD, "samplers.pt"))
    torch.save(q_sample_vectors(ref), os.path.join(GOLD, "q_sample.pt"))
    torch.save(unet_vectors(ref), os.path.join(GOLD, "unet_tiny.pt"))
    torch.save(unet3d_vectors(ref), os.path.join(GOLD, "unet3d_tiny.pt"))
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
