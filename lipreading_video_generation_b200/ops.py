"""Thin tensor-level wrappers over the C ABI (include/b2d.h).

Every function enqueues sm_100a kernels on torch's current CUDA stream through
ctypes; torch only owns the memory.  Activations are pixel-major fp16 tensors
(N, H, W, C) fp16 (gradient tensors carry a power-of-two loss scale) -- see DESIGN.md "data layout".
No function here has a CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L

F16 = torch.float16       # every 16-bit tensor (activations, packed weights, qkv, gradients) is fp16, fp32 accumulate
_F16_MAX = 65504.0
GRAD_SCALE_TARGET = 128.0  # loss scale 2^k puts max|d loss / d pred| * 2^k in [64, 128)


def _s():
    return L.stream_ptr()


class KernelTimer:
    """Optional CUDA-event timing of selected ABI calls on the launching stream (bench.py's roofline leg).
    `classes` is a set of names ("attention_fwd", "attention_bwd", "conv2d", "conv2d_wgrad");
    records (name, meta, start_event, end_event)."""
    active = None

    def __init__(self, classes):
        self.classes, self.records = set(classes), []

    def __enter__(self):
        KernelTimer.active = self
        return self

    def __exit__(self, *a):
        KernelTimer.active = None

    def totals(self):
        out = {}
        for name, meta, e0, e1 in self.records:
            d = out.setdefault((name, meta), [0.0, 0])
            d[0] += e0.elapsed_time(e1)
            d[1] += 1
        return out


class _timed:
    def __init__(self, name, meta=None):
        t = KernelTimer.active
        self.on = t is not None and name in t.classes
        self.name, self.meta = name, meta

    def __enter__(self):
        if self.on:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *a):
        if self.on:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            KernelTimer.active.records.append((self.name, self.meta, self.e0, e1))


pack_records = None      # a list while the engine records how one cached operand copy was made (see repack_all)
_PACK_DESC = np.dtype([("w", "<u8"), ("out", "<u8"), ("s_o", "<i8"), ("s_i", "<i8"), ("s_k", "<i8"), ("Cout", "<i4"),
                       ("Cin", "<i4"), ("K", "<i4"), ("transpose", "<i4"), ("rows_pad", "<i4"), ("inner_pad", "<i4"),
                       ("block_begin", "<i4"), ("reserved", "<i4")])      # B2dPackDesc, include/b2d.h


def _pack_on_device(w4: torch.Tensor, transpose: int, rows_pad: int, inner_pad: int) -> torch.Tensor:
    """One launch of b2d_pack_conv_weight on the (possibly strided) OIHW view; same values as the torch chain below."""
    co, ci, kh, kw = w4.shape
    refreshable = True
    if kh * kw > 1 and w4.stride(2) != kw * w4.stride(3):
        w4 = w4.contiguous()
        refreshable = False                 # packed from a temporary: cannot be re-packed from the parameter in place
    out = torch.empty(rows_pad, kh * kw * inner_pad, device=w4.device, dtype=F16)
    L.check(L.lib().b2d_pack_conv_weight(w4.data_ptr(), w4.stride(0), w4.stride(1), w4.stride(3), co, ci, kh * kw,
                                         transpose, out.data_ptr(), rows_pad, inner_pad, _s()), "b2d_pack_conv_weight")
    if pack_records is not None:
        pack_records.append((w4.data_ptr(), out, w4.stride(0), w4.stride(1), w4.stride(3), co, ci, kh * kw, transpose,
                             rows_pad, inner_pad) if refreshable else None)
    return out


def build_pack_table(records, block_elems: int):
    """Host logic of b2d_pack_conv_weights_multi: (B2dPackDesc[] as a numpy structured array, total CTAs)."""
    t = np.zeros(len(records), dtype=_PACK_DESC)
    blk = 0
    for i, (wptr, out, s_o, s_i, s_k, co, ci, K, tr, rp, ip) in enumerate(records):
        t[i] = (wptr, out if isinstance(out, int) else out.data_ptr(), s_o, s_i, s_k, co, ci, K, tr, rp, ip, blk, 0)
        blk += (rp * K * ip + block_elems - 1) // block_elems
    return t, blk


def repack_all(table_dev: torch.Tensor, n_descs: int, total_blocks: int) -> None:
    """Refresh every recorded fp16 operand copy from its fp32 parameter, in place, in ONE launch."""
    L.check(L.lib().b2d_pack_conv_weights_multi(table_dev.data_ptr(), n_descs, total_blocks, _s()),
            "b2d_pack_conv_weights_multi")


def pack_conv_weight(w: torch.Tensor, cin_pad: int | None = None, cout_pad_to: int = 16) -> torch.Tensor:
    """OIHW (or OI / OIK for Linear / Conv1d) fp32 -> fp16 [CoutPad][kh*kw*CinPad] (tap-major, channel-minor)."""
    if w.dim() == 2:
        w = w[:, :, None, None]
    elif w.dim() == 3:
        w = w[:, :, :, None]
    co, ci, kh, kw = w.shape
    cip = ci if cin_pad is None else cin_pad
    cop = (co + cout_pad_to - 1) // cout_pad_to * cout_pad_to
    if w.is_cuda and w.dtype == torch.float32:
        return _pack_on_device(w, 0, cop, cip)
    p = torch.zeros(cop, kh * kw, cip, device=w.device, dtype=F16)
    p[:co, :, :ci] = w.permute(0, 2, 3, 1).reshape(co, kh * kw, ci).clamp(-_F16_MAX, _F16_MAX).to(F16)
    return p.reshape(cop, kh * kw * cip)


def pack_dgrad_weight(w: torch.Tensor, cin_pad: int | None = None, cout_pad: int | None = None) -> torch.Tensor:
    """Weights of the data-gradient convolution: dX = conv(dY, flip(w)^T).
    OIHW -> fp16 [CinPad][kh*kw*CoutPad] with the taps spatially flipped."""
    if w.dim() == 2:
        w = w[:, :, None, None]
    elif w.dim() == 3:
        w = w[:, :, :, None]
    co, ci, kh, kw = w.shape
    cip = ci if cin_pad is None else cin_pad
    cop = co if cout_pad is None else cout_pad
    if w.is_cuda and w.dtype == torch.float32:
        return _pack_on_device(w, 1, (cip + 15) // 16 * 16, cop)
    wf = torch.flip(w, dims=(2, 3)).permute(1, 2, 3, 0).reshape(ci, kh * kw, co)  # [ci][tap'][co]
    p = torch.zeros((cip + 15) // 16 * 16, kh * kw, cop, device=w.device, dtype=F16)
    p[:ci, :, :co] = wf.clamp(-_F16_MAX, _F16_MAX).to(F16)
    return p.reshape(p.shape[0], kh * kw * cop)


def gn_fusable(Ho, Wo, cout) -> bool:
    """True when a convolution with this output plane / channel count can carry the fused GroupNorm epilogues."""
    return bool(L.lib().b2d_conv2d_gn_fusable(int(Ho), int(Wo), int(cout)))


class GnBwdFuse:
    """Arguments of the GroupNorm-backward prologue fused into a data-gradient convolution (b2d_conv_args.gnb_*):
    the convolution's output becomes dyh = dy * act'(z) * mask / (1 - p) and `ab` (fp32 (N, C, 2), zeroed by the
    caller) receives (sum dyh, sum dyh * x) per (sample, channel)."""
    __slots__ = ("x1", "x2", "mean", "rstd", "gamma", "beta", "act", "mask", "keep_scale", "ab")

    def __init__(self, x1, x2, mean, rstd, gamma, beta, act, ab, mask=None, keep_scale=1.0):
        self.x1, self.x2, self.mean, self.rstd, self.gamma, self.beta = x1, x2, mean, rstd, gamma, beta
        self.act, self.ab, self.mask, self.keep_scale = act, ab, mask, keep_scale


def conv2d(x, w_packed, cout, ksize=1, stride=1, x2=None, bias=None, emb=None, residual=None, out=None,
           out_nchw_f32=None, c1=None, c2=None, stats_out=None, gnb=None):
    """x: (N,H,W,ldx) fp16, optional x2 concatenated on channels.  Returns (N,Ho,Wo,cout) fp16 (or `out`).
    stats_out: fp64 (N, cout/2, 2) zeroed buffer that receives the per-(image, channel pair) sum / sum of squares of
    the output (consumed by groupnorm_apply); gnb: GnBwdFuse."""
    L.require_cuda(x)
    if any(t is not None and t.dtype != F16 for t in (x, x2, residual, w_packed)):
        raise RuntimeError("conv2d: x / x2 / residual / packed weights must be fp16")
    N, H, W, ldx = x.shape
    a = L.ConvArgs()
    a.x, a.ldx, a.C1 = x.data_ptr(), x.stride(2), (ldx if c1 is None else c1)
    if x2 is not None:
        a.x2, a.ldx2, a.C2 = x2.data_ptr(), x2.stride(2), (x2.shape[3] if c2 is None else c2)
    else:
        a.x2, a.ldx2, a.C2 = None, 0, 0
    a.N, a.H, a.W, a.ksize, a.stride = N, H, W, ksize, stride
    a.w, a.Cout, a.CoutPad = w_packed.data_ptr(), cout, w_packed.shape[0]
    a.bias, a.emb, a.ldemb = L.ptr(bias), L.ptr(emb), (emb.stride(0) if emb is not None else 0)
    a.residual, a.ldres = L.ptr(residual), (residual.stride(2) if residual is not None else 0)
    Ho, Wo = H // stride, W // stride
    if out is None and out_nchw_f32 is None:
        out = torch.empty(N, Ho, Wo, cout, device=x.device, dtype=x.dtype)
    a.out, a.ldo = L.ptr(out), (out.stride(2) if out is not None else 0)
    a.out_nchw_f32 = L.ptr(out_nchw_f32)
    a.gn_stats_out = L.ptr(stats_out)
    if gnb is not None:
        if gnb.ab.dtype != torch.float32 or gnb.x1.dtype != F16:
            raise RuntimeError("conv2d: gnb.ab must be fp32, gnb.x1 fp16")
        a.gnb_x1, a.gnb_ld1, a.gnb_C1 = gnb.x1.data_ptr(), gnb.x1.stride(2), gnb.x1.shape[3]
        a.gnb_x2, a.gnb_ld2 = (gnb.x2.data_ptr(), gnb.x2.stride(2)) if gnb.x2 is not None else (None, 0)
        a.gnb_mean, a.gnb_rstd = gnb.mean.data_ptr(), gnb.rstd.data_ptr()
        a.gnb_gamma, a.gnb_beta, a.gnb_act = gnb.gamma.data_ptr(), gnb.beta.data_ptr(), int(gnb.act)
        a.gnb_mask, a.gnb_keep_scale, a.gnb_ab = L.ptr(gnb.mask), float(gnb.keep_scale), gnb.ab.data_ptr()
    epi = "gnb" if gnb is not None else ("stats" if stats_out is not None else "")
    with _timed("conv2d", (N, H, W, a.C1 + a.C2, cout, ksize, stride) + ((epi,) if epi else ())):
        L.check(L.lib().b2d_conv2d(C.byref(a), _s()), "b2d_conv2d")
    return out if out is not None else out_nchw_f32


def conv2d_wgrad(x, dy, dw, ksize=1, stride=1, x2=None, c1=None, c2=None, cout=None):
    """Accumulate dw (fp32 [Cout][taps*(C1+C2)]) from x (+x2) and dy (N,Ho,Wo,Cout), all fp16."""
    if any(t is not None and t.dtype != F16 for t in (x, x2, dy)):
        raise RuntimeError("conv2d_wgrad: x / x2 / dy must be fp16")
    N, H, W, ldx = x.shape
    a = L.WgradArgs()
    a.x, a.ldx, a.C1 = x.data_ptr(), x.stride(2), (ldx if c1 is None else c1)
    if x2 is not None:
        a.x2, a.ldx2, a.C2 = x2.data_ptr(), x2.stride(2), (x2.shape[3] if c2 is None else c2)
    else:
        a.x2, a.ldx2, a.C2 = None, 0, 0
    a.N, a.H, a.W, a.ksize, a.stride = N, H, W, ksize, stride
    cout = dy.shape[3] if cout is None else cout
    ktot = ksize * ksize * (a.C1 + a.C2)
    co0 = 0
    while co0 < cout:  # kernel handles Cout == 64 or a multiple of 128 per launch
        rem = cout - co0
        chunk = rem - rem % 128 if rem >= 128 else 64
        a.dy, a.lddy, a.Cout = dy.data_ptr() + 2 * co0, dy.stride(2), chunk
        a.dw = dw.data_ptr() + 4 * co0 * ktot
        with _timed("conv2d_wgrad", (N, H, W, a.C1 + a.C2, chunk, ksize, stride)):
            L.check(L.lib().b2d_conv2d_wgrad(C.byref(a), _s()), "b2d_conv2d_wgrad")
        co0 += chunk
    return dw


_gn_ws: dict = {}


def _gn_workspace(dev, N, Ct):
    nbytes = L.lib().b2d_groupnorm_workspace_bytes(N, Ct)
    key = (dev, torch.cuda.current_stream().cuda_stream)
    ws = _gn_ws.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), device=dev, dtype=torch.uint8)
        _gn_ws[key] = ws
    return ws


def _gn_args(x1, x2, gamma, beta, eps, act, drop_p, seed, overflow=None):
    N = x1.shape[0]
    HW = x1.shape[1] * x1.shape[2]
    a = L.GnArgs()
    a.x1, a.ld1, a.C1 = x1.data_ptr(), x1.stride(2), x1.shape[3]
    if x2 is not None:
        a.x2, a.ld2, a.C2 = x2.data_ptr(), x2.stride(2), x2.shape[3]
    else:
        a.x2, a.ld2, a.C2 = None, 0, 0
    a.N, a.HW = N, HW
    a.gamma, a.beta, a.eps, a.act, a.drop_p, a.seed = gamma.data_ptr(), beta.data_ptr(), eps, int(act), drop_p, seed
    a.overflow = L.ptr(overflow)
    return a


def groupnorm_fwd(x1, gamma, beta, x2=None, act=True, eps=1e-5, drop_p=0.0, seed=0):
    """Returns (y fp16 (N,H,W,C1+C2), mean (N,32), rstd (N,32))."""
    L.require_cuda(x1)
    if x1.dtype != F16 or (x2 is not None and x2.dtype != F16):
        raise RuntimeError("groupnorm_fwd: inputs must be fp16")
    N, H, W, C1 = x1.shape
    Ct = C1 + (x2.shape[3] if x2 is not None else 0)
    a = _gn_args(x1, x2, gamma, beta, eps, act, drop_p, seed)
    y = torch.empty(N, H, W, Ct, device=x1.device, dtype=F16)
    mean = torch.empty(N, 32, device=x1.device, dtype=torch.float32)
    rstd = torch.empty(N, 32, device=x1.device, dtype=torch.float32)
    ws = _gn_workspace(x1.device, N, Ct)
    with _timed("groupnorm_fwd", (N, H * W, Ct)):
        L.check(L.lib().b2d_groupnorm_fwd(C.byref(a), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), ws.data_ptr(),
                                          ws.numel(), _s()), "b2d_groupnorm_fwd")
    return y, mean, rstd


def groupnorm_bwd(x1, gamma, beta, dy, mean, rstd, dgamma, dbeta, x2=None, act=True, eps=1e-5, drop_p=0.0, seed=0,
                  add_a=None, add_b=None, add_a_full=False, colsum_out=None, overflow=None):
    """Returns (dx1, dx2); dgamma/dbeta accumulated; add_a/add_b (fp16, any channel pitch) are added to dx1;
    colsum_out (fp32 (N, >= C1) view, accumulated): per-sample pixel sums of dx1."""
    N, H, W, C1 = x1.shape
    if any(t is not None and t.dtype != F16 for t in (x1, x2, dy, add_a, add_b)):
        raise RuntimeError("groupnorm_bwd: x / dy / add_a / add_b must be fp16")
    Ct = C1 + (x2.shape[3] if x2 is not None else 0)
    a = _gn_args(x1, x2, gamma, beta, eps, act, drop_p, seed, overflow)
    dx1 = torch.empty(N, H, W, C1, device=x1.device, dtype=F16)
    dx2 = torch.empty(N, H, W, x2.shape[3], device=x1.device, dtype=F16) if x2 is not None else None
    ws = _gn_workspace(x1.device, N, Ct)
    with _timed("groupnorm_bwd", (N, H * W, Ct)):
        L.check(L.lib().b2d_groupnorm_bwd(
            C.byref(a), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx1.data_ptr(), C1,
            L.ptr(dx2), (dx2.shape[3] if dx2 is not None else 0),
            L.ptr(add_a), (add_a.stride(2) if add_a is not None else 0),
            L.ptr(add_b), (add_b.stride(2) if add_b is not None else 0), int(add_a_full),
            L.ptr(colsum_out), (colsum_out.stride(0) if colsum_out is not None else 0),
            dgamma.data_ptr(), dbeta.data_ptr(), ws.data_ptr(), ws.numel(), _s()), "b2d_groupnorm_bwd")
    return dx1, dx2


def pair_stats(x, out=None):
    """fp64 (N, C/2, 2): per-(sample, channel pair) sum and sum of squares over the pixels of x (N,H,W,C) fp16 -- what a
    convolution with `stats_out` emits for its own output."""
    N, H, W, Cc = x.shape
    if out is None:
        out = torch.zeros(N, Cc // 2, 2, device=x.device, dtype=torch.float64)
    L.check(L.lib().b2d_pair_stats(x.data_ptr(), x.stride(2), Cc, N, H * W, out.data_ptr(), _s()), "b2d_pair_stats")
    return out


def groupnorm_apply(x1, st1, gamma, beta, x2=None, st2=None, act=True, eps=1e-5, drop_p=0.0, seed=0, rows_per_sample=1,
                    want_mask=False):
    """One-pass GroupNorm32 (+SiLU, +dropout) with the statistics taken from the producers' pair sums.
    Returns (y, mean, rstd, mask) -- mask (N, H, W, C/8) uint8 keep bits when want_mask and drop_p > 0, else None."""
    L.require_cuda(x1)
    if x1.dtype != F16 or (x2 is not None and x2.dtype != F16):
        raise RuntimeError("groupnorm_apply: inputs must be fp16")
    N, H, W, C1 = x1.shape
    Ct = C1 + (x2.shape[3] if x2 is not None else 0)
    a = _gn_args(x1, x2, gamma, beta, eps, act, drop_p, seed)
    y = torch.empty(N, H, W, Ct, device=x1.device, dtype=F16)
    mean = torch.empty(N, 32, device=x1.device, dtype=torch.float32)
    rstd = torch.empty(N, 32, device=x1.device, dtype=torch.float32)
    mask = torch.empty(N, H, W, Ct // 8, device=x1.device, dtype=torch.uint8) if (want_mask and drop_p > 0.0) else None
    with _timed("groupnorm_fwd", (N, H * W, Ct)):
        L.check(L.lib().b2d_groupnorm_apply(C.byref(a), st1.data_ptr(), L.ptr(st2), int(rows_per_sample), y.data_ptr(),
                                            mean.data_ptr(), rstd.data_ptr(), L.ptr(mask), _s()), "b2d_groupnorm_apply")
    return y, mean, rstd, mask


def groupnorm_bwd_apply(x1, gamma, beta, dyh, mean, rstd, ab, dgamma, dbeta, x2=None, eps=1e-5, add_a=None, add_b=None,
                        add_a_full=False, colsum_out=None, overflow=None):
    """Input-gradient pass of the GroupNorm backward given dyh / ab from the fused data-gradient epilogue."""
    N, H, W, C1 = x1.shape
    if any(t is not None and t.dtype != F16 for t in (x1, x2, dyh, add_a, add_b)):
        raise RuntimeError("groupnorm_bwd_apply: x / dyh / add_a / add_b must be fp16")
    a = _gn_args(x1, x2, gamma, beta, eps, True, 0.0, 0, overflow)
    dx1 = torch.empty(N, H, W, C1, device=x1.device, dtype=F16)
    dx2 = torch.empty(N, H, W, x2.shape[3], device=x1.device, dtype=F16) if x2 is not None else None
    Ct = C1 + (x2.shape[3] if x2 is not None else 0)
    with _timed("groupnorm_bwd", (N, H * W, Ct)):
        L.check(L.lib().b2d_groupnorm_bwd_apply(
            C.byref(a), dyh.data_ptr(), mean.data_ptr(), rstd.data_ptr(), ab.data_ptr(), dx1.data_ptr(), C1,
            L.ptr(dx2), (dx2.shape[3] if dx2 is not None else 0),
            L.ptr(add_a), (add_a.stride(2) if add_a is not None else 0),
            L.ptr(add_b), (add_b.stride(2) if add_b is not None else 0), int(add_a_full),
            L.ptr(colsum_out), (colsum_out.stride(0) if colsum_out is not None else 0),
            dgamma.data_ptr(), dbeta.data_ptr(), _s()), "b2d_groupnorm_bwd_apply")
    return dx1, dx2


_SM_COUNT: dict = {}


def _sm_count(dev):
    i = dev.index if dev.index is not None else torch.cuda.current_device()
    if i not in _SM_COUNT:
        _SM_COUNT[i] = torch.cuda.get_device_properties(i).multi_processor_count
    return _SM_COUNT[i]


def attention_fwd(qkv, heads, new_order=False, rows=None, split=None):
    """qkv (N,H,W,3C) fp16 -> (out (N,H,W,C) fp16, lse (N,heads,L) fp32).  rows = (q_row0, q_rows): sequence-parallel
    shard -- only those query rows of out / lse are written (the caller all-gathers `out`).  A grid that would leave
    most SMs idle (batch-1 sampling) takes the split-KV variant unless split=False."""
    if qkv.dtype != F16:
        raise RuntimeError("attention_fwd: qkv must be fp16")
    N, H, W, C3 = qkv.shape
    Cc = C3 // 3
    ch = Cc // heads
    out = torch.empty(N, H, W, Cc, device=qkv.device, dtype=F16)
    lse = torch.empty(N, heads, H * W, device=qkv.device, dtype=torch.float32)
    if rows is not None:
        with _timed("attention_fwd", (N, H * W, heads, ch, rows[1])):
            L.check(L.lib().b2d_attention_fwd_rows(qkv.data_ptr(), qkv.stride(2), N, H * W, heads, ch, int(new_order),
                                                   int(rows[0]), int(rows[1]), out.data_ptr(), Cc, lse.data_ptr(), _s()),
                    "b2d_attention_fwd_rows")
        return out, lse
    Lk = H * W
    ctas = N * heads * ((Lk + 255) // 256)
    sms = _sm_count(qkv.device)
    if ch in (64, 128) and Lk > 128 and 2 * ctas <= sms and split is not False:
        # small grid (batch-1 sampling): split the keys so that ~all SMs get a CTA
        splits = max(1, min(8, sms // ctas, ((Lk + (127 if ch == 64 else 63)) // (128 if ch == 64 else 64)) // 2))
        if splits > 1:
            nws = int(L.lib().b2d_attention_fwd_splitkv_ws_floats(N, Lk, heads, ch, splits))
            ws = torch.empty(nws, device=qkv.device, dtype=torch.float32)
            with _timed("attention_fwd", (N, Lk, heads, ch)):
                L.check(L.lib().b2d_attention_fwd_splitkv(qkv.data_ptr(), qkv.stride(2), N, Lk, heads, ch, int(new_order),
                                                          splits, ws.data_ptr(), nws, out.data_ptr(), Cc, lse.data_ptr(),
                                                          _s()), "b2d_attention_fwd_splitkv")
            return out, lse
    with _timed("attention_fwd", (N, H * W, heads, ch)):
        L.check(L.lib().b2d_attention_fwd(qkv.data_ptr(), qkv.stride(2), N, H * W, heads, ch, int(new_order),
                                          out.data_ptr(), Cc, lse.data_ptr(), _s()), "b2d_attention_fwd")
    return out, lse


def attention_bwd(qkv, out, d_out, lse, heads, new_order=False, rows=None):
    """qkv, out, d_out fp16 -> dqkv (N,H,W,3C) fp16.  rows = (q_row0, q_rows): this shard's partial dqkv (sum the
    shards' tensors for the full gradient)."""
    if qkv.dtype != F16 or out.dtype != F16 or d_out.dtype != F16:
        raise RuntimeError("attention_bwd: qkv / out / d_out must be fp16")
    N, H, W, C3 = qkv.shape
    Cc = C3 // 3
    ch = Cc // heads
    dqkv = torch.empty(N, H, W, C3, device=qkv.device, dtype=F16)
    delta = torch.empty(int(L.lib().b2d_attention_bwd_ws_floats(N, H * W, heads, ch)), device=qkv.device,
                        dtype=torch.float32)
    if rows is not None:
        with _timed("attention_bwd", (N, H * W, heads, ch, rows[1])):
            L.check(L.lib().b2d_attention_bwd_rows(qkv.data_ptr(), qkv.stride(2), out.data_ptr(), out.stride(2),
                                                   d_out.data_ptr(), d_out.stride(2), lse.data_ptr(), N, H * W, heads, ch,
                                                   int(new_order), int(rows[0]), int(rows[1]), dqkv.data_ptr(), C3,
                                                   delta.data_ptr(), _s()), "b2d_attention_bwd_rows")
        return dqkv
    with _timed("attention_bwd", (N, H * W, heads, ch)):
        L.check(L.lib().b2d_attention_bwd(qkv.data_ptr(), qkv.stride(2), out.data_ptr(), out.stride(2),
                                          d_out.data_ptr(), d_out.stride(2), lse.data_ptr(), N, H * W, heads, ch,
                                          int(new_order), dqkv.data_ptr(), C3, delta.data_ptr(), _s()),
                "b2d_attention_bwd")
    return dqkv


def assemble_input(xt, cond, wcond, audio, cpad):
    N, Cin, S, _ = xt.shape
    Cim = wcond.shape[0] if wcond is not None else 0
    Cc = cond.shape[1] if cond is not None else 0
    Hc, Wc = (cond.shape[2], cond.shape[3]) if cond is not None else (0, 0)
    A = audio.shape[1] if audio is not None else 0
    out = torch.empty(N, S, S, cpad, device=xt.device, dtype=F16)
    L.check(L.lib().b2d_assemble_input(xt.data_ptr(), L.ptr(cond), L.ptr(wcond), L.ptr(audio), N, S, Cin, Cc, Hc, Wc,
                                       Cim, A, cpad, out.data_ptr(), _s()), "b2d_assemble_input")
    return out


def assemble_input_bwd(dX, cond, cin, cim, A, d_wcond, d_audio):
    N, S, _, cpad = dX.shape
    Cc = cond.shape[1] if cond is not None else 0
    Hc, Wc = (cond.shape[2], cond.shape[3]) if cond is not None else (0, 0)
    L.check(L.lib().b2d_assemble_input_bwd(dX.data_ptr(), L.ptr(cond), N, S, cin, Cc, Hc, Wc, cim, A, cpad,
                                           L.ptr(d_wcond), L.ptr(d_audio), _s()), "b2d_assemble_input_bwd")


def upsample2x(x):
    N, H, W, Cc = x.shape
    y = torch.empty(N, 2 * H, 2 * W, Cc, device=x.device, dtype=x.dtype)       # pure 16-bit copy
    L.check(L.lib().b2d_upsample2x(x.data_ptr(), y.data_ptr(), N, H, W, Cc, _s()), "b2d_upsample2x")
    return y


def upsample2x_bwd(dy):
    N, H2, W2, Cc = dy.shape
    dx = torch.empty(N, H2 // 2, W2 // 2, Cc, device=dy.device, dtype=F16)
    L.check(L.lib().b2d_upsample2x_bwd(dy.data_ptr(), dx.data_ptr(), N, H2 // 2, W2 // 2, Cc, _s()), "b2d_upsample2x_bwd")
    return dx


def zero_insert2x(dy):
    N, Ho, Wo, Cc = dy.shape
    z = torch.empty(N, 2 * Ho, 2 * Wo, Cc, device=dy.device, dtype=dy.dtype)     # pure 16-bit copy
    L.check(L.lib().b2d_zero_insert2x(dy.data_ptr(), z.data_ptr(), N, Ho, Wo, Cc, _s()), "b2d_zero_insert2x")
    return z


def nchw_to_pixel_major(src, cpad, scale=None):
    """fp32 NCHW -> pixel-major fp16, multiplied by the device scalar scale[0] when given (loss scale)."""
    N, Cc, H, W = src.shape
    dst = torch.empty(N, H, W, cpad, device=src.device, dtype=F16)
    L.check(L.lib().b2d_nchw_to_pixel_major(src.data_ptr(), dst.data_ptr(), N, Cc, H * W, cpad, L.ptr(scale), _s()),
            "b2d_nchw_to_pixel_major")
    return dst


def grad_scale(x, target=GRAD_SCALE_TARGET, state=None, reduce_max=None):
    """Device-chosen power-of-two loss scale for the fp16 gradient tensors: returns fp32 [scale, 1/scale].
    state: int32[4] overflow / back-off state (b2d_grad_scale_finalize); reduce_max: callable applied to the 1-element
    max|x| tensor between the two halves -- ranks that sum 16-bit partial gradients must agree on the scale."""
    s2 = torch.empty(2, device=x.device, dtype=torch.float32)
    L.check(L.lib().b2d_absmax(x.data_ptr(), x.numel(), s2.data_ptr(), _s()), "b2d_absmax")
    if reduce_max is not None:
        reduce_max(s2[:1])
    L.check(L.lib().b2d_grad_scale_finalize(s2.data_ptr(), float(target), L.ptr(state), _s()), "b2d_grad_scale_finalize")
    return s2


def scale_f32_(x, scale_ptr_tensor):
    """x *= scale_ptr_tensor[0] in place (x fp32 contiguous view)."""
    L.check(L.lib().b2d_scale_f32(x.data_ptr(), x.numel(), scale_ptr_tensor.data_ptr(), _s()), "b2d_scale_f32")
    return x


def colsum(x, out, c_off=0, C_=None):
    """out[n][c] += sum over pixels of x[n,:,:,c_off+c];  x (N,H,W,ld) fp16, out (N,C) fp32."""
    N, H, W, ld = x.shape
    C_ = out.shape[1] if C_ is None else C_
    L.check(L.lib().b2d_colsum(x.data_ptr(), x.stride(2), c_off, C_, N, H * W, out.data_ptr(), out.stride(0), _s()), "b2d_colsum")
    return out


_W2V_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def layernorm_fwd(x2, gamma, beta, eps):
    """nn.LayerNorm over the last dim of x2 (R, C) fp32 / bf16 / fp16 -> (y fp32 (R, C), mean (R,), rstd (R,))."""
    L.require_cuda(x2)
    R, Cc = x2.shape
    y = torch.empty(R, Cc, device=x2.device, dtype=torch.float32)
    mean = torch.empty(R, device=x2.device, dtype=torch.float32)
    rstd = torch.empty(R, device=x2.device, dtype=torch.float32)
    L.check(L.lib().b2d_layernorm_fwd(x2.data_ptr(), _W2V_DT[x2.dtype], gamma.data_ptr(), beta.data_ptr(), R, Cc, eps,
                                      y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), _s()), "b2d_layernorm_fwd")
    return y, mean, rstd


def layernorm_bwd(x2, gamma, mean, rstd, dy):
    """-> (dx fp32 (R, C), dgamma (C,), dbeta (C,))."""
    R, Cc = x2.shape
    dx = torch.empty(R, Cc, device=x2.device, dtype=torch.float32)
    dgb = torch.zeros(2, Cc, device=x2.device, dtype=torch.float32)
    L.check(L.lib().b2d_layernorm_bwd(x2.data_ptr(), _W2V_DT[x2.dtype], gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                      dy.data_ptr(), R, Cc, dx.data_ptr(), dgb[0].data_ptr(), dgb[1].data_ptr(), _s()),
            "b2d_layernorm_bwd")
    return dx, dgb[0], dgb[1]


def colsum_rows16(x2):
    """fp32 (C,) column sums of a bf16 / fp16 (R, C) matrix (last dim contiguous)."""
    L.require_cuda(x2)
    R, Cc = x2.shape
    out = torch.empty(Cc, device=x2.device, dtype=torch.float32)
    L.check(L.lib().b2d_colsum_rows16(x2.data_ptr(), _W2V_DT[x2.dtype], x2.stride(0), R, Cc, out.data_ptr(), _s()),
            "b2d_colsum_rows16")
    return out


def transpose16(x):
    """(B, R, C) contiguous bf16 / fp16 -> (B, C, R) contiguous."""
    L.require_cuda(x)
    B, R, Cc = x.shape
    out = torch.empty(B, Cc, R, device=x.device, dtype=x.dtype)
    L.check(L.lib().b2d_transpose16(x.data_ptr(), out.data_ptr(), B, R, Cc, _s()), "b2d_transpose16")
    return out


def w2v_layer0_fwd(audio, w, bias, gamma, beta, stride, eps, out_dtype):
    """Fused first feature-extractor layer of wav2vec2 (conv k=10 -> per-channel GroupNorm -> erf GELU).
    audio (N,S) fp32, w (C,1,K) fp32 -> out (N,C,T) in out_dtype, mean (N,C), rstd (N,C)."""
    N, S = audio.shape
    Cc, _, K = w.shape
    T = (S - K) // stride + 1
    out = torch.empty(N, Cc, T, device=audio.device, dtype=out_dtype)
    mean = torch.empty(N, Cc, device=audio.device, dtype=torch.float32)
    rstd = torch.empty_like(mean)
    L.check(L.lib().b2d_w2v_layer0_fwd(audio.data_ptr(), audio.stride(0), w.data_ptr(),
                                       None if bias is None else bias.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                       N, S, Cc, K, stride, float(eps), out.data_ptr(), _W2V_DT[out_dtype],
                                       mean.data_ptr(), rstd.data_ptr(), _s()), "b2d_w2v_layer0_fwd")
    return out, mean, rstd


def w2v_layer0_bwd(audio, w, bias, gamma, beta, stride, eps, dout, mean, rstd):
    """Parameter gradients of the fused layer: (dw (C,1,K), dbias (C) or None, dgamma (C), dbeta (C)), fp32."""
    N, S = audio.shape
    Cc, _, K = w.shape
    flat = torch.zeros(Cc * (K + 3), device=audio.device, dtype=torch.float32)
    dw, dbias, dgamma, dbeta = flat[:Cc * K], flat[Cc * K:Cc * (K + 1)], flat[Cc * (K + 1):Cc * (K + 2)], flat[Cc * (K + 2):]
    L.check(L.lib().b2d_w2v_layer0_bwd(audio.data_ptr(), audio.stride(0), w.data_ptr(),
                                       None if bias is None else bias.data_ptr(), gamma.data_ptr(), beta.data_ptr(),
                                       N, S, Cc, K, stride, float(eps), dout.data_ptr(), _W2V_DT[dout.dtype],
                                       mean.data_ptr(), rstd.data_ptr(), dw.data_ptr(),
                                       None if bias is None else dbias.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(),
                                       _s()), "b2d_w2v_layer0_bwd")
    return dw.view(Cc, 1, K), (None if bias is None else dbias), dgamma, dbeta


def _sa_strides(q, k, v):
    B, H, Lq, D = q.shape
    if not (q.shape == k.shape == v.shape and q.dtype == k.dtype == v.dtype and q.stride() == k.stride() == v.stride()
            and q.stride(3) == 1):
        raise ValueError("short_attention: q, k, v must share shape, dtype and strides with a unit head-dim stride")
    return B, H, Lq, D, q.stride(0), q.stride(2), q.stride(1)


def short_attention_fwd(q, k, v, scale, drop_p=0.0, seed=0):
    """q, k, v (B, H, L, 64) strided views, L <= 16 -> out (B, L, H, 64) contiguous."""
    B, H, Lq, D, sb, sl, sh = _sa_strides(q, k, v)
    out = torch.empty(B, Lq, H, D, device=q.device, dtype=q.dtype)
    L.check(L.lib().b2d_short_attention_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), sb, sl, sh, B, H, Lq, D,
                                            _W2V_DT[q.dtype], float(scale), float(drop_p), int(seed), out.data_ptr(), _s()),
            "b2d_short_attention_fwd")
    return out


def short_attention_bwd(q, k, v, dout, scale, drop_p=0.0, seed=0):
    """Gradients (dq, dk, dv), each (B, L, H, 64) contiguous; dout (B, L, H, 64) contiguous."""
    B, H, Lq, D, sb, sl, sh = _sa_strides(q, k, v)
    g = torch.empty(3, B, Lq, H, D, device=q.device, dtype=q.dtype)
    L.check(L.lib().b2d_short_attention_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), sb, sl, sh, dout.data_ptr(), B, H, Lq,
                                            D, _W2V_DT[q.dtype], float(scale), float(drop_p), int(seed), g[0].data_ptr(),
                                            g[1].data_ptr(), g[2].data_ptr(), _s()), "b2d_short_attention_bwd")
    return g[0], g[1], g[2]


def reduce_rows(inp, out):
    R, Cc = inp.shape
    L.check(L.lib().b2d_reduce_rows(inp.data_ptr(), R, Cc, inp.stride(0), out.data_ptr(), _s()), "b2d_reduce_rows")
    return out


_REDUCE_DESC = np.dtype([("in", "<u8"), ("out", "<u8"), ("ld", "<i8"), ("C", "<i4"), ("block_begin", "<i4")])


def reduce_rows_multi(jobs):
    """out[c] += sum_r inp[r][c] for every (inp (R, C) fp32, out (C,)) in `jobs` (one R) with ONE launch."""
    if not jobs:
        return
    R = jobs[0][0].shape[0]
    tab = np.zeros(len(jobs), dtype=_REDUCE_DESC)
    blocks = 0
    for i, (inp, out) in enumerate(jobs):
        if inp.shape[0] != R or inp.dtype != torch.float32 or out.dtype != torch.float32 or inp.stride(1) != 1:
            raise RuntimeError("reduce_rows_multi: jobs must be fp32 (R, C) row-major views with one R")
        tab[i] = (inp.data_ptr(), out.data_ptr(), inp.stride(0), inp.shape[1], blocks)
        blocks += (inp.shape[1] + 31) // 32
    dev = jobs[0][0].device
    pinned = torch.from_numpy(tab.view("uint8").reshape(-1)).pin_memory()
    table = pinned.to(dev, non_blocking=True)
    L.check(L.lib().b2d_reduce_rows_multi(table.data_ptr(), len(jobs), R, blocks, _s()), "b2d_reduce_rows_multi")
    table.record_stream(torch.cuda.current_stream(dev))


def timestep_embedding(t, dim, freqs):
    N = t.shape[0]
    out = torch.empty(N, dim, device=t.device, dtype=torch.float32)
    L.check(L.lib().b2d_timestep_embedding(t.data_ptr(), N, dim, freqs.data_ptr(), out.data_ptr(), _s()),
            "b2d_timestep_embedding")
    return out


def linear_fwd(x, w, bias, silu_in=False, relu_out=False):
    M, K = x.shape
    Nout = w.shape[0]
    y = torch.empty(M, Nout, device=x.device, dtype=torch.float32)
    L.check(L.lib().b2d_linear_fwd(x.data_ptr(), w.data_ptr(), L.ptr(bias), y.data_ptr(), M, K, Nout, int(silu_in),
                                   int(relu_out), _s()), "b2d_linear_fwd")
    return y


def linear_bwd(x, w, y, dy, dw, db, need_dx=True, silu_in=False, relu_out=False):
    M, K = x.shape
    Nout = w.shape[0]
    dx = torch.empty(M, K, device=x.device, dtype=torch.float32) if need_dx else None
    L.check(L.lib().b2d_linear_bwd(x.data_ptr(), w.data_ptr(), L.ptr(y), dy.data_ptr(), L.ptr(dx), dw.data_ptr(),
                                   L.ptr(db), M, K, Nout, int(silu_in), int(relu_out), _s()), "b2d_linear_bwd")
    return dx
