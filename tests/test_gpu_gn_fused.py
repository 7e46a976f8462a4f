"""GPU parity of the GroupNorm pieces fused into the convolution epilogues (north_star item 2: "fused
GroupNorm/SiLU/residual epilogues"; GroupNorm32 utils.py:54-56 at unet.py:194-198,218-225,297,624-628):
  * b2d_conv2d with gn_stats_out: per-(image, channel pair) moments of the output,
  * b2d_groupnorm_apply: one-pass GroupNorm(+SiLU)(+dropout) from those moments, incl. a two-tensor skip concatenation
    whose groups straddle the two tensors,
  * b2d_conv2d with gnb_*: the data-gradient convolution that folds act'(z) and the dropout mask into its output and
    emits the channel sums, followed by b2d_groupnorm_bwd_apply,
each against torch's fp32 GroupNorm / autograd on the CPU-side values (the oracle's group_norm32 is F.group_norm)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def gen(seed):
    return torch.Generator().manual_seed(seed)


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def _nchw(t):                         # pixel-major (N,H,W,C) -> the reference's NCHW, fp32 on the CPU
    return t.detach().float().cpu().permute(0, 3, 1, 2).contiguous()


@pytest.mark.parametrize("N,H,W,cin,cout,k,extras", [
    (3, 32, 32, 64, 128, 3, False), (5, 8, 8, 128, 64, 1, True), (2, 16, 16, 64, 256, 3, True),
    (2, 16, 16, 128, 192, 1, False), (1, 64, 64, 64, 64, 3, True)])
@pytest.mark.parametrize("halo", ["0", "1"])
def test_conv_epilogue_pair_statistics(N, H, W, cin, cout, k, extras, halo, monkeypatch):
    """gn_stats_out of a convolution == pair sums of its own output (one-pass kernel and torch), incl. tiles that hold
    several images with a ragged last tile (N = 5 at 8x8), two N tiles (256 channels), bias + emb + residual.
    halo = "1": the opt-in haloed-patch variant of the 3x3 kernel (conv_gemm.cu) where the shape allows it."""
    from lipreading_video_generation_b200 import ops
    monkeypatch.setenv("B2D_CONV_HALO", halo)
    g = gen(N * H + cout)
    x = torch.randn(N, H, W, cin, generator=g).half().cuda()
    w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).cuda()
    bias = torch.randn(cout, generator=g).cuda() if extras else None
    emb = torch.randn(N, cout, generator=g).cuda() if extras else None
    res = torch.randn(N, H, W, cout, generator=g).half().cuda() if extras else None
    assert ops.gn_fusable(H, W, cout)
    st = torch.zeros(N, cout // 2, 2, device="cuda", dtype=torch.float64)
    out = ops.conv2d(x, ops.pack_conv_weight(w), cout, k, 1, bias=bias, emb=emb, residual=res, stats_out=st)
    plain = ops.conv2d(x, ops.pack_conv_weight(w), cout, k, 1, bias=bias, emb=emb, residual=res)
    assert torch.equal(out, plain), "the statistics epilogue must not change the output"
    o = out.double().view(N, H * W, cout // 2, 2)
    want = torch.stack([o.sum(dim=(1, 3)), (o * o).sum(dim=(1, 3))], dim=-1)          # (N, C/2, 2)
    rms = o.pow(2).mean().sqrt().item()
    n_el = 2 * H * W
    # the epilogue sums the fp32 values before the fp16 rounding of the store: differences are rounding noise
    assert (st[..., 0] - want[..., 0]).abs().max().item() <= 2e-3 * rms * n_el ** 0.5 + 1e-6
    assert rel(st[..., 1], want[..., 1]) <= 1e-3
    one_pass = ops.pair_stats(out)
    assert rel(one_pass, want) <= 1e-6


@pytest.mark.parametrize("N,H,W,C1,C2,act", [(2, 32, 32, 64, 0, True), (3, 16, 16, 256, 128, True),
                                             (2, 16, 16, 128, 64, False), (2, 8, 8, 256, 256, True)])
def test_groupnorm_apply_from_pair_statistics(N, H, W, C1, C2, act):
    """One-pass GroupNorm32 (+SiLU) with moments from pair sums == F.group_norm on the concatenation; (256 | 128)
    channels puts group 21 across the two tensors (12 channels per group)."""
    from lipreading_video_generation_b200 import ops
    g = gen(C1 + C2 + H)
    x1 = (torch.randn(N, H, W, C1, generator=g) * 1.5 + 0.3).half().cuda()
    x2 = (torch.randn(N, H, W, C2, generator=g) * 0.7 - 0.2).half().cuda() if C2 else None
    Ct = C1 + C2
    gamma = (1 + 0.1 * torch.randn(Ct, generator=g)).cuda()
    beta = (0.1 * torch.randn(Ct, generator=g)).cuda()
    st1 = ops.pair_stats(x1)
    st2 = ops.pair_stats(x2) if C2 else None
    y, mean, rstd, mask = ops.groupnorm_apply(x1, st1, gamma, beta, x2=x2, st2=st2, act=act)
    assert mask is None
    xc = torch.cat([_nchw(x1)] + ([_nchw(x2)] if C2 else []), dim=1)
    ref = F.group_norm(xc, 32, gamma.cpu(), beta.cpu(), 1e-5)
    ref = F.silu(ref) if act else ref
    assert rel(_nchw(y), ref) <= 2e-3
    xg = xc.view(N, 32, -1)
    assert torch.allclose(mean.cpu(), xg.mean(dim=2), atol=1e-5, rtol=1e-5)
    assert torch.allclose(rstd.cpu(), 1.0 / torch.sqrt(xg.var(dim=2, unbiased=False) + 1e-5), rtol=1e-4)
    # the two-kernel path gives the same values
    y_old, mean_old, rstd_old = ops.groupnorm_fwd(x1, gamma, beta, x2=x2, act=act)
    assert rel(y, y_old) <= 1e-3 and torch.allclose(mean, mean_old, atol=1e-5)


def test_groupnorm_apply_rows_per_sample_folds_frames():
    """dims=3: a sample is T frames that were convolved as separate images -- the statistics rows of its frames are
    summed (rows_per_sample = T)."""
    from lipreading_video_generation_b200 import ops
    g = gen(5)
    B, T, H, W, C = 2, 3, 8, 8, 64
    x = torch.randn(B * T, H, W, C, generator=g).half().cuda()
    gamma, beta = torch.ones(C).cuda(), torch.zeros(C).cuda()
    st = ops.pair_stats(x)                                   # per frame
    folded = x.view(B, T * H, W, C)
    y, mean, rstd, _ = ops.groupnorm_apply(folded, st, gamma, beta, act=False, rows_per_sample=T)
    xc = _nchw(folded)
    assert rel(_nchw(y), F.group_norm(xc, 32, None, None, 1e-5)) <= 2e-3


def test_groupnorm_apply_dropout_mask_bits_match_the_output():
    from lipreading_video_generation_b200 import ops
    g = gen(6)
    x = torch.randn(2, 16, 16, 128, generator=g).half().cuda()
    gamma, beta = torch.ones(128).cuda(), torch.zeros(128).cuda()
    st = ops.pair_stats(x)
    y0, _, _, _ = ops.groupnorm_apply(x, st, gamma, beta, act=True)
    y1, _, _, mask = ops.groupnorm_apply(x, st, gamma, beta, act=True, drop_p=0.1, seed=77, want_mask=True)
    assert mask.shape == (2, 16, 16, 16) and mask.dtype == torch.uint8
    bits = ((mask.unsqueeze(-1) >> torch.arange(8, device="cuda", dtype=torch.uint8)) & 1).view(2, 16, 16, 128).bool()
    assert abs(bits.float().mean().item() - 0.9) < 0.01
    assert torch.equal(y1[~bits], torch.zeros_like(y1[~bits]))
    assert torch.allclose(y1[bits].float(), y0[bits].float() / 0.9, rtol=2e-3, atol=1e-4)


@pytest.mark.parametrize("N,H,W,C1,C2,cnext,k,act,drop", [
    (2, 32, 32, 64, 0, 64, 3, True, 0.0), (2, 16, 16, 128, 64, 128, 3, True, 0.0), (3, 16, 16, 256, 0, 768, 1, False, 0.0),
    (2, 32, 32, 128, 0, 128, 3, True, 0.1), (5, 8, 8, 64, 0, 64, 3, True, 0.1), (2, 16, 16, 256, 128, 256, 3, True, 0.0)])
@pytest.mark.parametrize("halo", ["0", "1"])
def test_groupnorm_backward_fused_into_the_data_gradient_convolution(N, H, W, C1, C2, cnext, k, act, drop, halo, monkeypatch):
    """y = dropout(act(GroupNorm([x1 | x2]))), o = conv(y): given dL/do, the data-gradient convolution with the gnb
    epilogue + b2d_groupnorm_bwd_apply must equal autograd of the same chain (fp32, same dropout mask): dx1, dx2 (with
    the fused residual adds), dgamma, dbeta, and the column sums of dx1.  halo = "1": haloed-patch variant (3x3 only)."""
    from lipreading_video_generation_b200 import ops
    monkeypatch.setenv("B2D_CONV_HALO", halo)
    g = gen(C1 + C2 + cnext + H)
    Ct = C1 + C2
    x1 = (torch.randn(N, H, W, C1, generator=g) + 0.2).half().cuda()
    x2 = (torch.randn(N, H, W, C2, generator=g) * 0.8).half().cuda() if C2 else None
    gamma = (1 + 0.1 * torch.randn(Ct, generator=g)).cuda()
    beta = (0.1 * torch.randn(Ct, generator=g)).cuda()
    w = (torch.randn(cnext, Ct, k, k, generator=g) / (Ct * k * k) ** 0.5)
    d_o = torch.randn(N, H, W, cnext, generator=g).half().cuda()
    add_a = torch.randn(N, H, W, C1, generator=g).half().cuda()
    st1 = ops.pair_stats(x1)
    st2 = ops.pair_stats(x2) if C2 else None
    y, mean, rstd, mask = ops.groupnorm_apply(x1, st1, gamma, beta, x2=x2, st2=st2, act=act, drop_p=drop, seed=5,
                                              want_mask=True)
    ab = torch.zeros(N, Ct, 2, device="cuda")
    fuse = ops.GnBwdFuse(x1, x2, mean, rstd, gamma, beta, act, ab, mask=mask, keep_scale=1.0 / (1.0 - drop) if drop else 1.0)
    dyh = ops.conv2d(d_o, ops.pack_dgrad_weight(w.cuda()), Ct, k, 1, gnb=fuse)
    dgamma, dbeta = torch.zeros(Ct, device="cuda"), torch.zeros(Ct, device="cuda")
    colsum = torch.zeros(N, C1, device="cuda")
    ovf = torch.zeros(1, device="cuda", dtype=torch.int32)
    dx1, dx2 = ops.groupnorm_bwd_apply(x1, gamma, beta, dyh, mean, rstd, ab, dgamma, dbeta, x2=x2, add_a=add_a,
                                       colsum_out=colsum, overflow=ovf)
    # ---- reference: autograd on the CPU in fp32 with the same mask
    xs = [_nchw(x1).requires_grad_(True)] + ([_nchw(x2).requires_grad_(True)] if C2 else [])
    gm, bt = gamma.cpu().clone().requires_grad_(True), beta.cpu().clone().requires_grad_(True)
    z = F.group_norm(torch.cat(xs, dim=1), 32, gm, bt, 1e-5)
    a = F.silu(z) if act else z
    if drop:
        bits = ((mask.unsqueeze(-1) >> torch.arange(8, device="cuda", dtype=torch.uint8)) & 1).view(N, H, W, Ct)
        a = a * _nchw(bits) / (1.0 - drop)
    o = F.conv2d(a, w, padding=k // 2)
    o.backward(_nchw(d_o))
    assert rel(_nchw(dx1), xs[0].grad + _nchw(add_a)) <= 5e-3
    if C2:
        assert rel(_nchw(dx2), xs[1].grad) <= 5e-3
    assert rel(dgamma, gm.grad) <= 5e-3 and rel(dbeta, bt.grad) <= 5e-3
    assert rel(colsum, _nchw(dx1).sum(dim=(2, 3))) <= 2e-3
    assert int(ovf.item()) == 0
    # the unfused two-kernel backward computes the same thing from the raw dy
    dy = ops.conv2d(d_o, ops.pack_dgrad_weight(w.cuda()), Ct, k, 1)
    dg2, db2 = torch.zeros(Ct, device="cuda"), torch.zeros(Ct, device="cuda")
    dx1_old, _ = ops.groupnorm_bwd(x1, gamma, beta, dy, mean, rstd, dg2, db2, x2=x2, act=act, drop_p=drop, seed=5, add_a=add_a)
    assert rel(dx1, dx1_old) <= 5e-3


def test_groupnorm_backward_overflow_counter_flags_saturated_gradients():
    """fp16_util.py:236-237 check_overflow: a gradient tensor that saturated at +-65504 must be REPORTED (device
    counter, no host sync in the step), not silently clamped."""
    from lipreading_video_generation_b200 import ops
    g = gen(9)
    x = torch.randn(2, 16, 16, 64, generator=g).half().cuda()
    gamma, beta = torch.ones(64).cuda(), torch.zeros(64).cuda()
    st = ops.pair_stats(x)
    y, mean, rstd, _ = ops.groupnorm_apply(x, st, gamma, beta, act=True)
    dyh = torch.randn(2, 16, 16, 64, generator=g).half().cuda()
    dyh[0, 3, 3, 5] = 65504.0
    ab = torch.zeros(2, 64, 2, device="cuda")
    ovf = torch.zeros(1, device="cuda", dtype=torch.int32)
    ops.groupnorm_bwd_apply(x, gamma, beta, dyh, mean, rstd, ab, torch.zeros(64).cuda(), torch.zeros(64).cuda(), overflow=ovf)
    assert int(ovf.item()) >= 1


def test_loss_scale_backs_off_after_an_overflowed_step_and_adam_skips_it():
    """The device-side form of MixedPrecisionTrainer (fp16_util.py:148-233): the saturation counter of one backward
    (state[0]) makes the optimiser step a no-op (b2d_adam_step_guarded) and lowers the next backward's loss scale by one
    bit (b2d_grad_scale_finalize); clean steps raise it again by 1/1024 bit each.  No host synchronisation involved."""
    import lipreading_video_generation_b200 as P
    from lipreading_video_generation_b200 import ops
    x = (torch.randn(4096, generator=gen(3)) * 1e-3).cuda()
    clean = ops.grad_scale(x).tolist()
    state = torch.tensor([7, 0, 0, 0], dtype=torch.int32, device="cuda")
    after = ops.grad_scale(x, state=state).tolist()
    assert after[0] == clean[0] / 2 and after[1] == clean[1] * 2
    assert state.tolist() == [0, 1024, 1, 7]
    again = ops.grad_scale(x, state=state).tolist()                   # a clean step: 1023/1024 bit still rounds up to 1
    assert again[0] == clean[0] / 2 and state.tolist() == [0, 1023, 1, 0]
    state[1] = 2
    assert ops.grad_scale(x, state=state).tolist()[0] == clean[0] / 2 and state.tolist()[1] == 1
    assert ops.grad_scale(x, state=state).tolist()[0] == clean[0] and state.tolist()[1] == 0
    # guarded Adam
    w = torch.nn.Parameter(torch.ones(1000, device="cuda"))
    w.grad = torch.full((1000,), 0.5, device="cuda")
    flag = torch.zeros(4, dtype=torch.int32, device="cuda")
    opt = P.Adam([w], lr=1e-2, skip_on_overflow=flag)
    flag[0] = 3
    opt.step()
    assert torch.equal(w.detach(), torch.ones(1000, device="cuda")), "an overflowed step must leave the weights untouched"
    flag[0] = 0
    opt.step()
    assert (w.detach() < 1.0).all()
