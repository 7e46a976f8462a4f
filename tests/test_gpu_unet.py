"""GPU parity (through the C ABI): UNetAudio forward / loss / gradients vs the reference's golden vectors
and vs the CPU oracle.  Tolerance: north_star's bf16 bar, relative error <= 2e-2 (per-tensor ||a-b||/||b||),
applied to the output, the loss and EVERY parameter gradient."""
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


def build_tiny(g):
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    cfg = dict(g["cfg"])
    model = P.UNetAudio(**cfg, audio_encoder=_StubEncoder())
    model.load_state_dict(R.fill_params(model.state_dict(), seed=g["seed"]))
    return model.cuda().train()


def _tiny_step(golden):
    import lipreading_video_generation_b200 as P
    g = golden("unet_tiny.pt")
    model = build_tiny(g)
    sched = P.LinearNoiseScheduler(1000, 0.00085, 0.012)
    x0, eps, t = g["x0"].cuda(), g["eps"].cuda(), g["t"].cuda()
    xt = sched.add_noise(x0, eps, t)
    pred = model(xt, g["cond"].cuda(), {"feats": g["feats"].cuda()}, t)
    loss = P.MSELoss()(pred, eps)
    return g, model, xt, pred, loss


def test_tiny_forward_and_loss_match_reference_golden(golden):
    g, model, xt, pred, loss = _tiny_step(golden)
    assert torch.equal(xt.cpu(), g["xt"]), "q_sample must be bit-exact"
    assert pred.shape == g["pred"].shape and pred.dtype == torch.float32
    e = rel(pred, g["pred"])
    assert e <= BF16_TOL, f"denoiser output rel err {e}"
    assert abs(loss.item() - g["loss"].item()) <= BF16_TOL * abs(g["loss"].item())


def test_tiny_grad_norms_match_reference_golden(golden):
    """Structural check of the hand-scheduled backward: every parameter's gradient norm (all tensors)."""
    g, model, xt, pred, loss = _tiny_step(golden)
    loss.backward()
    named = dict(model.named_parameters())
    bad = []
    for k, n_ref in g["grad_norms"].items():
        n = named[k].grad.float().norm().item()
        if abs(n - n_ref) > 3 * BF16_TOL * n_ref + 1e-7:
            bad.append((k, n, n_ref))
    assert not bad, bad[:5]


# History: with bf16 storage this test missed the bar (worst 2.01e-2 tiny, 3.3e-2..5.6e-2 full config;
# profiles/parity_r1_status.md).  All 16-bit tensors are now fp16 (11-bit significand, same tensor-core rate) with a
# device-chosen power-of-two loss scale on the gradient tensors.  The tolerance was not changed.
def test_tiny_gradients_match_reference_golden_within_bf16_bar(golden):
    g, model, xt, pred, loss = _tiny_step(golden)
    loss.backward()
    named = dict(model.named_parameters())
    worst = ("", 0.0)
    for k, ref_g in g["grads"].items():
        e = rel(named[k].grad, ref_g)
        if e > worst[1]:
            worst = (k, e)
    assert worst[1] <= BF16_TOL, f"gradient {worst[0]} rel err {worst[1]}"


def test_eval_forward_no_grad_matches_train_forward(golden):
    g = golden("unet_tiny.pt")
    model = build_tiny(g)
    xt, t = g["xt"].cuda(), g["t"].cuda()
    a = model(xt, g["cond"].cuda(), {"feats": g["feats"].cuda()}, t)
    model.eval()
    with torch.no_grad():
        b = model(xt, g["cond"].cuda(), {"feats": g["feats"].cuda()}, t)
    # same kernels, same schedule; GroupNorm statistics are accumulated with fp32 shared-memory atomics, so two
    # runs may differ by bf16 re-rounding of a few activations (not a parity bar: both are checked vs golden)
    assert rel(a, b) <= 2e-3


def test_cpu_input_fails_loudly(golden):
    g = golden("unet_tiny.pt")
    model = build_tiny(g)
    with pytest.raises(RuntimeError):
        model(g["xt"], g["cond"], {"feats": g["feats"]}, g["t"])


def _full_config_step(S, N, init, seed=3):
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    cfg = dict(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
               attention_resolutions=(1, 2, 4), dropout=0.0, audio_feature_dim=768, projected_audio_dim=128)
    model = P.UNetAudio(**cfg, audio_encoder=_StubEncoder())
    sd = getattr(R, init)(model.state_dict(), seed=seed)
    model.load_state_dict(sd)
    model = model.cuda().train()
    g = torch.Generator().manual_seed(5)
    x0 = torch.rand(N, 3, S, S, generator=g) * 2 - 1
    cond = torch.rand(N, 3, S, S, generator=g) * 2 - 1
    feats = torch.randn(N, 12, 768, generator=g) * 0.5
    eps = torch.randn(N, 3, S, S, generator=g)
    t = torch.randint(0, 1000, (N,), generator=g)
    xt = R.add_noise(R.linear_tables(1000, 0.00085, 0.012), x0, eps, t)
    pred = model(xt.cuda(), cond.cuda(), {"feats": feats.cuda()}, t.cuda())
    loss = P.MSELoss()(pred, eps.cuda())
    loss.backward()
    sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    rcfg = dict(image_size=S, model_channels=64, num_res_blocks=2, attention_resolutions=(1, 2, 4),
                channel_mult=(1, 2, 4), num_heads=1)
    rp = R.unet_audio_forward(sdr, rcfg, xt, cond, feats.mean(1), t)
    rl = R.mse(rp, eps)
    rl.backward()
    errs = sorted((rel(p.grad, sdr[k].grad), k) for k, p in model.named_parameters())
    return rel(pred, rp), loss.item(), rl.item(), errs


@pytest.mark.parametrize("init", ["protocol_params", "fill_params"])
def test_train_py_config_64x64_output_loss_and_all_309_gradients_vs_oracle(init):
    """The exact train.py:88-97 model (ch 64, mult (1,2,4), 2 res blocks, attention at every level) at 64x64,
    BASELINE.md protocol init (torch-default + un_zero'd zero_modules) and the golden-fixture init."""
    e_pred, loss, loss_ref, errs = _full_config_step(64, 2, init)
    assert e_pred <= BF16_TOL, e_pred
    assert abs(loss - loss_ref) <= BF16_TOL * abs(loss_ref)
    assert len(errs) == 309
    assert errs[-1][0] <= BF16_TOL, f"worst gradient {errs[-1]}, median {errs[len(errs) // 2][0]}"


def test_sampler_trajectory_step_by_step_vs_oracle(golden):
    """test.py:57-65 loop: 6 DDPM steps with injected x_T and z; every x_t compared with the oracle trajectory
    (oracle model forward + oracle sampler update).  Also checks the V2 sampler test.py:111 uses, and DDIM."""
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    g = golden("unet_tiny.pt")
    model = build_tiny(g).eval()
    sd = R.fill_params(model.state_dict(), seed=g["seed"])
    c = g["cfg"]
    rcfg = dict(image_size=c["image_size"], model_channels=c["model_channels"], num_res_blocks=c["num_res_blocks"],
                attention_resolutions=c["attention_resolutions"], channel_mult=c["channel_mult"], num_heads=1)
    gg = torch.Generator().manual_seed(77)
    N, S, T = 2, c["image_size"], 6
    xT = torch.randn(N, 3, S, S, generator=gg)
    zs = [torch.randn(N, 3, S, S, generator=gg) for _ in range(T)]
    cond, feats = g["cond"], g["feats"]
    for kind in ("ddpm", "v2"):
        if kind == "ddpm":
            sched, tab, step = P.LinearNoiseScheduler(T, 0.00085, 0.012), R.linear_tables(T, 0.00085, 0.012), R.ddpm_step
        else:
            sched, tab, step = P.LinearNoiseSchedulerV2(T, 5e-5, 0.015), R.linear_tables(T, 5e-5, 0.015), R.ddpm_v2_step
        traj = []
        P.sample_images(model, sched, cond.cuda(), {"feats": feats.cuda()}, T, xt=xT.cuda(), noises=[z.cuda() for z in zs],
                        callback=lambda i, x, x0: traj.append((i, x.cpu(), x0.cpu())))
        x = xT
        with torch.no_grad():
            for k, i in enumerate(reversed(range(T))):
                e = R.unet_audio_forward(sd, rcfg, x, cond, feats.mean(1), torch.full((N,), i))
                x, x0 = step(tab, x, e, i, zs[k])
                assert traj[k][0] == i
                assert rel(traj[k][1], x) <= BF16_TOL and rel(traj[k][2], x0) <= BF16_TOL, (kind, i)
    xa, _ = P.sample_images(model, P.LinearNoiseScheduler(50, 0.00085, 0.012), cond.cuda(), {"feats": feats.cuda()}, 50,
                            xt=xT.cuda(), ddim_steps=4)
    assert torch.isfinite(xa).all() and xa.shape == xT.shape


# ------------------------------------------------------------------ dims=3
def _tiny3d_step(golden):
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    g = golden("unet3d_tiny.pt")
    model = P.UNetModel(**dict(g["cfg"]))
    model.load_state_dict(R.fill_params(model.state_dict(), seed=g["seed"]))
    model = model.cuda().train()
    pred = model(g["x"].cuda(), g["t"].cuda())
    loss = P.MSELoss()(pred, g["target"].cuda())
    return g, model, pred, loss


def test_unet3d_state_dict_keys_and_shapes_match_reference_golden(golden):
    import lipreading_video_generation_b200 as P
    g = golden("unet3d_tiny.pt")
    model = P.UNetModel(**dict(g["cfg"]))
    sd = model.state_dict()
    assert set(sd) == set(g["grad_norms"]), set(sd) ^ set(g["grad_norms"])
    for k, v in g["grads"].items():
        assert tuple(sd[k].shape) == tuple(v.shape), k


def test_unet3d_forward_loss_and_gradients_match_reference_golden(golden):
    """UNetModel(dims=3) on (B, C, T, S, S) = (2, 64, 4, 16, 16): Conv3d as three time-shifted 2-D tap groups on the
    frame-folded batch, GroupNorm / attention over T*S*S positions, Downsample (1,2,2), Upsample (T,2H,2W) -- against the
    real reference class (fixture made by oracle/make_golden.py), north_star's 16-bit bar 2e-2 on the output, the
    loss, the stored gradients and (x3, as a norm check) every gradient norm."""
    g, model, pred, loss = _tiny3d_step(golden)
    assert pred.shape == g["pred"].shape and pred.dtype == torch.float32
    e = rel(pred, g["pred"])
    assert e <= BF16_TOL, f"dims=3 output rel err {e}"
    assert abs(loss.item() - g["loss"].item()) <= BF16_TOL * abs(g["loss"].item())
    loss.backward()
    named = dict(model.named_parameters())
    worst = ("", 0.0)
    for k, ref_g in g["grads"].items():
        assert tuple(named[k].grad.shape) == tuple(ref_g.shape), k
        e = rel(named[k].grad, ref_g)
        if e > worst[1]:
            worst = (k, e)
    assert worst[1] <= BF16_TOL, worst
    bad = []
    for k, n_ref in g["grad_norms"].items():
        n = named[k].grad.float().norm().item()
        if abs(n - n_ref) > 3 * BF16_TOL * n_ref + 1e-7:
            bad.append((k, n, n_ref))
    assert not bad, bad[:5]


def test_unet3d_single_frame_equals_centre_taps(golden):
    """T = 1: only the centre temporal tap can touch a frame, so dims=3 on (B, C, 1, S, S) must equal the same
    network evaluated frame by frame with no temporal neighbours -- checked against the CPU oracle."""
    import lipreading_video_generation_b200 as P
    from oracle import restatement as R
    g = golden("unet3d_tiny.pt")
    model = P.UNetModel(**dict(g["cfg"]))
    sd = R.fill_params(model.state_dict(), seed=g["seed"])
    model.load_state_dict(sd)
    model = model.cuda().eval()
    x = g["x"][:, :, :1].contiguous()
    with torch.no_grad():
        pred = model(x.cuda(), g["t"].cuda())
        ref = R.unet_forward(sd, g["oracle_cfg"], x, g["t"])
    assert rel(pred, ref) <= BF16_TOL


def test_sampling_through_a_cuda_graph_equals_eager_sampling(golden):
    """GraphedDenoiser: the model forward recorded once and replayed per step gives the eager trajectory (same x_T, same
    injected z) at every step; one replay = the forward's kernels without their launch calls; update() re-records after
    a parameter change."""
    import lipreading_video_generation_b200 as P
    from lipreading_video_generation_b200 import _lib as L
    g = golden("unet_tiny.pt")
    model = build_tiny(g).eval()
    c = g["cfg"]
    gg = torch.Generator().manual_seed(78)
    N, S, T = 2, c["image_size"], 5
    xT = torch.randn(N, 3, S, S, generator=gg).cuda()
    zs = [torch.randn(N, 3, S, S, generator=gg).cuda() for _ in range(T)]
    cond, audio = g["cond"].cuda(), {"feats": g["feats"].cuda()}
    sched = P.LinearNoiseScheduler(T, 0.00085, 0.012)
    eager, graphed = [], []
    P.sample_images(model, sched, cond, audio, T, xt=xT, noises=zs, callback=lambda i, x, x0: eager.append(x.clone()))
    P.sample_images(model, sched, cond, audio, T, xt=xT, noises=zs, graph=True,
                    callback=lambda i, x, x0: graphed.append(x.clone()))
    assert len(eager) == len(graphed) == T
    for k in range(T):
        # the GroupNorm pair statistics are summed by fp64 atomics: run-to-run differences are rounding noise
        assert rel(graphed[k], eager[k]) <= 1e-3, (k, rel(graphed[k], eager[k]))
    # a replay launches nothing through the ABI; the recording did
    feats = model.encode_audio(audio)
    gd = P.GraphedDenoiser(model, cond, feats)
    assert gd.launches_per_replay > 20
    t = torch.full((N,), 3, device="cuda", dtype=torch.long)
    l0 = L.lib().b2d_launch_count()
    e_graph = gd(xT, 3).clone()
    assert L.lib().b2d_launch_count() == l0
    with torch.no_grad():
        e_eager = model(xT, cond, audio, t, audio_features=feats)
    assert rel(e_graph, e_eager) <= 5e-3
    # parameters change -> update() -> the replay follows the new weights
    with torch.no_grad():
        for p in model.out[2].parameters():
            p.mul_(1.5)
    gd.update()
    with torch.no_grad():
        e_eager2 = model(xT, cond, audio, t, audio_features=feats)
    assert rel(e_eager2, e_eager) > 0.1
    assert rel(gd(xT, 3), e_eager2) <= 5e-3
