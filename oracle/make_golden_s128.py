"""Generate the BASELINE configs[2]/[3]-shape fixtures (128x128, L = 16384) from the REAL reference (imported from
/root/reference) and pin oracle/restatement.py against it at those sizes.  Run in the build container:

    python oracle/make_golden_s128.py

Writes
  tests/golden/unet_s128.pt  -- the exact train.py:88-97 model at 128x128, N = 1: output, loss, the norm of every one of
                                the 309 torso gradients and a seeded sample of <= 4096 elements of each (full tensors
                                would be 65 MB); weights are regenerated from a seed (restatement.protocol_params).
  tests/golden/attn_big.pt   -- QKVAttentionLegacy (unet.py:349-366) forward + autograd at (L, ch) = (16384, 64) and
                                (4096, 128): 512 sampled positions of the output and of d qkv (inputs regenerated from
                                a seed), plus their full norms.
The GPU box only loads these files; nothing here runs there.
"""
from __future__ import annotations

import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_import, restatement as R  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
S128_CFG = dict(image_size=128, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
                attention_resolutions=(1, 2, 4), dropout=0.0, audio_feature_dim=768, projected_audio_dim=128)
SAMPLE = 4096


class _StubEncoder(torch.nn.Module):
    def forward(self, audio):
        return audio["feats"]


def s128_inputs(seed=7, N=1, S=128):
    g = torch.Generator().manual_seed(seed)
    x0 = torch.rand(N, 3, S, S, generator=g) * 2 - 1
    cond = torch.rand(N, 3, S, S, generator=g) * 2 - 1
    feats = torch.randn(N, 12, 768, generator=g) * 0.5
    eps = torch.randn(N, 3, S, S, generator=g)
    t = torch.randint(0, 1000, (N,), generator=g)
    return x0, cond, feats, eps, t


def sample_index(name: str, numel: int, k: int = SAMPLE):
    """Seeded element sample of a gradient tensor (the same indices are rebuilt in the GPU test)."""
    if numel <= k:
        return torch.arange(numel)
    seed = sum((i + 1) * ord(c) for i, c in enumerate(name)) % (2 ** 31)
    g = torch.Generator().manual_seed(seed)
    return torch.randperm(numel, generator=g)[:k]


def unet_s128(ref):
    torch.manual_seed(0)
    model = ref.unet_audio.UNetAudio(**S128_CFG)
    model.audio_encoder = _StubEncoder()
    sd = R.protocol_params({k: v for k, v in model.state_dict().items()}, seed=3)
    model.load_state_dict(sd)
    model.train()
    x0, cond, feats, eps, t = s128_inputs()
    xt = ref.linear_noise_scheduler.LinearNoiseScheduler(1000, 0.00085, 0.012).add_noise(x0, eps, t)
    t0 = time.time()
    pred = model(xt, cond, {"feats": feats}, t)
    loss = torch.nn.MSELoss()(pred, eps)
    loss.backward()
    print(f"reference fwd+bwd at 128x128, N=1: {time.time() - t0:.1f} s, loss {loss.item():.6f}")
    grads = {k: p.grad.detach() for k, p in model.named_parameters() if p.grad is not None}
    assert len(grads) == 309, len(grads)
    # ---- pin the restatement at this size: forward only (its autograd would keep five 1.07 GB softmax matrices x2)
    with torch.no_grad():
        rp = R.unet_audio_forward(sd, dict(image_size=128, model_channels=64, num_res_blocks=2,
                                           attention_resolutions=(1, 2, 4), channel_mult=(1, 2, 4), num_heads=1),
                                  xt, cond, feats.mean(dim=1), t)
    d = (rp - pred.detach()).abs().max().item()
    assert d <= 5e-5, d
    print(f"restatement pinned at 128x128: max|pred diff| {d:.2e}")
    out = dict(cfg=dict(S128_CFG), init="protocol_params", seed=3, input_seed=7, xt=xt, pred=pred.detach().clone(),
               loss=loss.detach().clone(), grad_norms={k: v.norm().item() for k, v in grads.items()},
               grad_samples={k: v.reshape(-1)[sample_index(k, v.numel())].clone() for k, v in grads.items()})
    return out


def attn_big(ref):
    out = {}
    for L, ch, seed in ((16384, 64, 41), (4096, 128, 42)):
        g = torch.Generator().manual_seed(seed)
        qkv = torch.randn(1, 3 * ch, L, generator=g).half().float()     # the reference's (B, 3C, L) layout; values
        d_o = torch.randn(1, ch, L, generator=g).half().float()         # exactly representable in the 16-bit storage type
        qkv.requires_grad_(True)
        t0 = time.time()
        a = ref.unet.QKVAttentionLegacy(1)(qkv)
        a.backward(d_o)
        print(f"reference QKVAttentionLegacy L={L} ch={ch}: {time.time() - t0:.1f} s")
        with torch.no_grad():
            ra = R.qkv_attention_legacy(qkv.detach(), 1)
        assert torch.equal(ra, a.detach()) or (ra - a.detach()).abs().max().item() <= 1e-6
        pos = torch.randperm(L, generator=torch.Generator().manual_seed(seed + 100))[:512].sort().values
        out[(L, ch)] = dict(seed=seed, pos=pos, out=a.detach()[0][:, pos].clone(), dqkv=qkv.grad[0][:, pos].clone(),
                            out_norm=a.detach().norm().item(), dqkv_norm=qkv.grad.norm().item())
    return out


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    ref = ref_import.load()
    os.makedirs(GOLD, exist_ok=True)
    if "--attn-only" not in sys.argv:
        torch.save(unet_s128(ref), os.path.join(GOLD, "unet_s128.pt"))
    torch.save(attn_big(ref), os.path.join(GOLD, "attn_big.pt"))
    for f in ("attn_big.pt", "unet_s128.pt"):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
