"""GPU parity at the BASELINE configs[2]/[3] shapes (128x128 images, L = 16384 attention positions) against fixtures
produced by the REAL reference (oracle/make_golden_s128.py; the GPU box only loads them), plus the data-parallel
gradient identity (SURVEY section 4 test 5) on one device.  Tolerances are north_star's 16-bit bar (2e-2, per tensor
||a-b||/||b||) end to end and the per-kernel bars of tests/test_gpu_kernels.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu

BF16_TOL = 2e-2


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


class _StubEncoder(torch.nn.Module):
    def forward(self, audio):
        return audio["feats"]


def _sample_index(name, numel, k=4096):
    """Same seeded element sample as oracle/make_golden_s128.sample_index."""
    if numel <= k:
        return torch.arange(numel)
    seed = sum((i + 1) * ord(c) for i, c in enumerate(name)) % (2 ** 31)
    return torch.randperm(numel, generator=torch.Generator().manual_seed(seed))[:k]


@pytest.mark.parametrize("L,ch,side", [(16384, 64, 128), (4096, 128, 64)])
def test_attention_at_full_resolution_lengths_vs_reference_golden(golden, L, ch, side):
    """QKVAttentionLegacy (unet.py:349-366) at the 128x128 model's full-resolution (L = 16384, ch 64: 128 key tiles per
    query tile) and half-resolution (L = 4096, ch 128) shapes: output and d qkv at 512 sampled positions, and the
    norms of the whole tensors, against the reference class run on the CPU."""
    from lipreading_video_generation_b200 import ops
    g = golden("attn_big.pt")[(L, ch)]
    gen = torch.Generator().manual_seed(g["seed"])
    qkv_ref = torch.randn(1, 3 * ch, L, generator=gen).half()
    d_o_ref = torch.randn(1, ch, L, generator=gen).half()
    qkv = qkv_ref.permute(0, 2, 1).contiguous().view(1, side, side, 3 * ch).cuda()       # pixel-major (N, H, W, 3C)
    d_o = d_o_ref.permute(0, 2, 1).contiguous().view(1, side, side, ch).cuda()
    out, lse = ops.attention_fwd(qkv, 1, False)
    assert torch.isfinite(lse).all()
    o = out.view(L, ch).float().cpu()
    pos = g["pos"]
    assert rel(o[pos].t(), g["out"]) <= 5e-3
    assert abs(o.norm().item() - g["out_norm"]) <= 5e-3 * g["out_norm"]
    dqkv = ops.attention_bwd(qkv, out, d_o, lse, 1, False).view(L, 3 * ch).float().cpu()
    assert torch.isfinite(dqkv).all()
    assert rel(dqkv[pos].t(), g["dqkv"]) <= 1e-2
    assert abs(dqkv.norm().item() - g["dqkv_norm"]) <= 1e-2 * g["dqkv_norm"]


def test_train_py_config_128x128_output_loss_and_all_309_gradients_vs_reference_golden(golden):
    """The exact train.py:88-97 model at 128x128 (BASELINE configs[2]/[3] image size; attention over 16384 / 4096 /
    1024 positions), N = 1: output, loss, the norm of EVERY parameter gradient and a 4096-element sample of each
    against the real reference's values."""
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    g = golden("unet_s128.pt")
    model = P.UNetAudio(**dict(g["cfg"]), audio_encoder=_StubEncoder())
    model.load_state_dict(getattr(R, g["init"])(model.state_dict(), seed=g["seed"]))
    model = model.cuda().train()
    gi = torch.Generator().manual_seed(g["input_seed"])
    S = g["cfg"]["image_size"]
    _x0 = torch.rand(1, 3, S, S, generator=gi)
    cond = torch.rand(1, 3, S, S, generator=gi) * 2 - 1
    feats = torch.randn(1, 12, 768, generator=gi) * 0.5
    eps = torch.randn(1, 3, S, S, generator=gi)
    t = torch.randint(0, 1000, (1,), generator=gi)
    pred = model(g["xt"].cuda(), cond.cuda(), {"feats": feats.cuda()}, t.cuda())
    e = rel(pred, g["pred"])
    assert e <= BF16_TOL, f"output rel err {e}"
    loss = P.MSELoss()(pred, eps.cuda())
    assert abs(loss.item() - g["loss"].item()) <= BF16_TOL * abs(g["loss"].item())
    loss.backward()
    named = dict(model.named_parameters())
    assert len(g["grad_samples"]) == 309
    errs = []
    for k, ref_s in g["grad_samples"].items():
        gr = named[k].grad.detach().float().cpu().reshape(-1)
        errs.append((rel(gr[_sample_index(k, gr.numel())], ref_s), k))
        n_ref = g["grad_norms"][k]
        assert abs(gr.norm().item() - n_ref) <= 3 * BF16_TOL * n_ref + 1e-9, (k, gr.norm().item(), n_ref)
    errs.sort()
    assert errs[-1][0] <= BF16_TOL, f"worst gradient {errs[-1]}, median {errs[len(errs) // 2][0]}"


def _dp_model(seed=3):
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    cfg = dict(image_size=32, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=1,
               attention_resolutions=(1, 2), channel_mult=(1, 2), dropout=0.0, audio_feature_dim=768,
               projected_audio_dim=128)
    model = P.UNetAudio(**cfg, audio_encoder=_StubEncoder())
    model.load_state_dict(R.protocol_params(model.state_dict(), seed=seed))
    return model.cuda().train()


def test_data_parallel_average_of_shard_gradients_equals_full_batch_gradient():
    """SURVEY section 4 test 5 on one device: the mean of the per-shard gradients (what `GradientAllReducer`'s AVG
    all-reduce produces; each rank's loss is the mean over its own clips, train.py:132) equals the gradient of the
    concatenated batch.  The 2-rank NCCL run of the same identity is scripts/dp_check.py (profiles/r2_dp_check.md)."""
    import lipreading_video_generation_b200 as P
    model = _dp_model()
    g = torch.Generator().manual_seed(17)
    N, S = 4, 32
    xt = torch.randn(N, 3, S, S, generator=g).cuda()
    cond = (torch.rand(N, 3, S, S, generator=g) * 2 - 1).cuda()
    feats = (torch.randn(N, 12, 768, generator=g) * 0.5).cuda()
    eps = torch.randn(N, 3, S, S, generator=g).cuda()
    t = torch.randint(0, 1000, (N,), generator=g).cuda()
    crit = P.MSELoss()

    def grads(lo, hi):
        model.zero_grad(set_to_none=True)
        crit(model(xt[lo:hi], cond[lo:hi], {"feats": feats[lo:hi]}, t[lo:hi]), eps[lo:hi]).backward()
        return {k: p.grad.detach().float().clone() for k, p in model.named_parameters() if p.grad is not None}
    full = grads(0, N)
    a, b = grads(0, N // 2), grads(N // 2, N)
    worst = max((rel((a[k] + b[k]) / 2, full[k]), k) for k in full)
    assert worst[0] <= 5e-3, worst          # same kernels, different batch tiling: 16-bit re-rounding only


def test_gradient_all_reducer_single_rank_nccl_keeps_gradients_and_fills_missing_ones():
    """The reducer on a real (1-rank) NCCL communicator: torso buckets on the side stream, the non-torso parameters
    through the persistent flat buffer with a FIXED layout -- a parameter whose layer was skipped this step (wav2vec2
    LayerDrop) contributes zeros instead of changing the bucket length (round-1 advisor finding)."""
    import os
    import torch.distributed as dist
    import lipreading_video_generation_b200 as P
    from lipreading_video_generation_b200.parallel import GradientAllReducer
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29541")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))

    class Enc(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.used = torch.nn.Parameter(torch.ones(768))
            self.skipped = torch.nn.Parameter(torch.ones(5))

        def forward(self, a):
            return a["feats"] * self.used

    model = _dp_model()
    model.audio_encoder = Enc().cuda()
    g = torch.Generator().manual_seed(18)
    N, S = 2, 32
    xt = torch.randn(N, 3, S, S, generator=g).cuda()
    cond = (torch.rand(N, 3, S, S, generator=g) * 2 - 1).cuda()
    feats = (torch.randn(N, 12, 768, generator=g) * 0.5).cuda()
    eps = torch.randn(N, 3, S, S, generator=g).cuda()
    t = torch.randint(0, 1000, (N,), generator=g).cuda()
    crit = P.MSELoss()
    crit(model(xt, cond, {"feats": feats}, t), eps).backward()
    want = {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}
    model.zero_grad(set_to_none=True)
    red = GradientAllReducer(model, bucket_mb=0.25, force=True)
    for _ in range(2):                               # second step: persistent buffers are reused
        model.zero_grad(set_to_none=True)
        crit(model(xt, cond, {"feats": feats}, t), eps).backward()
        red.finish()
        torch.cuda.synchronize()
        got = {k: p.grad for k, p in model.named_parameters()}
        for k, w in want.items():
            assert rel(got[k], w) <= 5e-3, k      # two runs of one step differ by 16-bit re-rounding (atomic order in the sums)
        assert got["audio_encoder.skipped"] is not None and float(got["audio_encoder.skipped"].abs().sum()) == 0.0
    assert red.bytes_reduced > 0
    dist.destroy_process_group()


def test_wav2vec2_bf16_autocast_encoder_is_inside_the_16bit_bar_of_the_fp32_stock_encoder():
    """bench.py runs the third-party HF wav2vec2 encoder (unet_audio.py:10-18) under bf16 autocast with three pieces
    replaced by this library's kernels; the reference runs the stock module in fp32 (train.py:97).  At the bench
    shape (256 windows of 4000 samples, eval mode so that LayerDrop / dropout / SpecAugment draw nothing): last hidden
    state and the time-mean the denoiser consumes, against the untouched fp32 HF module with the same weights."""
    import copy
    from transformers import Wav2Vec2Config, Wav2Vec2Model
    import lipreading_video_generation_b200.unet_audio as UA
    torch.manual_seed(0)
    enc = UA.Wav2Vec2Encoder(pretrained=False)
    stock = Wav2Vec2Model(Wav2Vec2Config())
    stock.load_state_dict(copy.deepcopy(enc.wav2vec2.state_dict()))
    enc, stock = enc.cuda().eval(), stock.cuda().eval()
    audio = torch.randn(256, 4000, generator=torch.Generator().manual_seed(5)).cuda()
    with torch.no_grad():
        ref = stock(input_values=audio).last_hidden_state
        with torch.autocast("cuda", dtype=torch.bfloat16):
            got = enc({"input_values": audio})
    assert got.shape == ref.shape == (256, 12, 768)
    assert rel(got, ref) <= BF16_TOL, rel(got, ref)
    assert rel(got.float().mean(dim=1), ref.mean(dim=1)) <= BF16_TOL


def test_deferred_weight_gradients_equal_the_in_order_backward():
    """The engine enqueues the last part of the torso's weight gradients on a side stream after the torso backward, so
    that they overlap the audio encoder's backward, and hands the gradients over in an end-of-backward callback.  The
    result must equal the in-order backward (first step: nothing is deferred because the work total is unknown)."""
    import lipreading_video_generation_b200 as P

    class Enc(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.scale = torch.nn.Parameter(torch.ones(768))

        def forward(self, a):
            return a["feats"] * self.scale

    model = _dp_model()
    model.audio_encoder = Enc().cuda()
    eng = model.engine()
    g = torch.Generator().manual_seed(21)
    N, S = 2, 32
    xt = torch.randn(N, 3, S, S, generator=g).cuda()
    cond = (torch.rand(N, 3, S, S, generator=g) * 2 - 1).cuda()
    feats = (torch.randn(N, 12, 768, generator=g) * 0.5).cuda()
    eps = torch.randn(N, 3, S, S, generator=g).cuda()
    t = torch.randint(0, 1000, (N,), generator=g).cuda()
    crit = P.MSELoss()

    def step():
        model.zero_grad(set_to_none=True)
        crit(model(xt, cond, {"feats": feats}, t), eps).backward()
        torch.cuda.synchronize()
        return {k: p.grad.detach().float().clone() for k, p in model.named_parameters()}
    assert eng.defer_frac > 0.0 and eng._wg_total == 0.0
    first = step()                                  # in order
    assert eng._wg_total > 0.0
    second = step()                                 # deferred tail
    eng.defer_frac = 0.0
    third = step()                                  # in order again
    for k in first:
        assert rel(second[k], first[k]) <= 5e-3, k
        assert rel(third[k], first[k]) <= 5e-3, k
    assert set(second) == set(first) and all(v is not None for v in second.values())
