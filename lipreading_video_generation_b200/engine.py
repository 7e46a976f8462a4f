"""DenoiserEngine -- hand-scheduled forward / backward of the UNet torso over the C-ABI kernels.

Replaces autograd through unet.py:646-675 (UNetModel.forward) and unet_audio.py:53-66 (conditioning
assembly) with a static layer schedule: every arithmetic step is one call into libb200diff.so;
PyTorch owns memory, streams and the parameter objects only.  The whole torso is ONE
torch.autograd.Function (`denoiser_apply`), so the reference's training recipe
(`loss.backward(); optimizer.step()`, train.py:131-134) works unchanged.

Layout: activations and activation gradients pixel-major (N,H,W,C) fp16 (gradients carry a power-of-two loss
scale chosen on the device each step; fp32 parameter gradients are unscaled stage by stage); skip concatenations (unet.py:672) are never
materialised (two-source kernels); gradient accumulation for residual / skip fan-out is fused into the
GroupNorm-backward and dgrad epilogues.
"""
from __future__ import annotations

import os

import torch

from . import ops
from .unet import AttentionBlock, Downsample, ResBlock, Upsample, timestep_freqs


def _pad64(c: int) -> int:
    return (c + 63) // 64 * 64


class _ZeroArena:
    """Zero-filled scratch for one forward (or one backward): the small accumulation buffers of a pass -- GroupNorm
    pair statistics, per-channel gradient sums, bias / embedding column sums -- are slices of ONE buffer cleared by ONE
    memset at the start of the pass instead of a `torch.zeros` (allocation + fill kernel) each.  The size is learnt
    on the first pass; a request that does not fit the cleared part falls back to `torch.zeros` and grows the next
    pass's buffer.  Slices are only valid until the next `begin()`: nothing handed out here is kept across passes."""

    def __init__(self):
        self.buf, self.off, self.need, self.cleared, self.device = None, 0, 0, 0, None

    def begin(self, device):
        self.device = device
        if self.need and (self.buf is None or self.buf.device != device or self.buf.numel() < self.need):
            self.buf = torch.empty(self.need, device=device, dtype=torch.uint8)
        self.cleared = min(self.need, self.buf.numel()) if self.buf is not None else 0
        if self.cleared:
            self.buf[:self.cleared].zero_()
        self.off = 0

    def zeros(self, shape, dtype):
        n = 1
        for d in shape:
            n *= d
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        lo = self.off
        self.off = (lo + nbytes + 255) // 256 * 256
        self.need = max(self.need, self.off)
        if self.off <= self.cleared:
            return self.buf[lo:lo + nbytes].view(dtype).view(shape)
        return torch.zeros(shape, device=self.device, dtype=dtype)


class _Ctx:
    """Saved tensors of one layer (forward -> backward)."""
    __slots__ = ("kind", "mod", "t")

    def __init__(self, kind, mod):
        self.kind, self.mod, self.t = kind, mod, {}


class DenoiserEngine:
    def __init__(self, model):
        self.model = model
        self.is_audio = hasattr(model, "audio_transformer")
        self.cin = model.in_channels                      # UNet input channels (195 for UNetAudio)
        self.cin_pad = _pad64(self.cin)
        self.resblocks = [m for m in model.modules() if isinstance(m, ResBlock)]
        self.emb_off = {}
        off = 0
        for rb in self.resblocks:
            self.emb_off[id(rb)] = off
            off += rb.out_channels
        self.emb_total = off
        self._freqs = None
        self._wcache = {}
        self._pack_table = None
        self._weights_maybe_stale = False
        self._colsums = {}
        self._stats = {}                      # data_ptr -> (pair statistics, tensor) of tensors a fused convolution produced
        self._fwd_arena, self._bwd_arena = _ZeroArena(), _ZeroArena()
        self._overflow = None                 # int32 device counter: saturated fp16 gradients seen by the backward
        # A/B switch for the fused epilogues: 0 = two-kernel GroupNorm, 1 = fused statistics + fused backward reduce,
        # 2 = fused statistics only (backward: two-kernel reduce + apply)
        mode = os.environ.get("B2D_GN_FUSE", "1")
        self.fuse_groupnorm = mode != "0"
        self.fuse_groupnorm_bwd = mode == "1"
        # The fused epilogues cost 30-90 us on the convolutions that are already bound by their epilogue or their L2
        # traffic (1x1 and 64-channel 3x3 layers: K = taps * Cin small, the tensor pipe 2 % busy); on the K >= 1152
        # layers they are nearly free (profiles/r2_groupnorm_fused.md).  With 8 epilogue warps per CTA the backward
        # epilogue (which removes the reduce pass AND the second SiLU' / dropout evaluation of the two-kernel path) wins
        # on every layer, the statistics epilogue only from K = 1152 on: below that the statistics come from one pass
        # over the tensor (b2d_pair_stats).  B2D_GN_FUSE_MINK="stats,gnb" overrides (0,0 = fuse everywhere).
        mk = os.environ.get("B2D_GN_FUSE_MINK", "1152,0").split(",")
        self.fuse_min_k_stats, self.fuse_min_k_gnb = int(mk[0]), int(mk[1])
        # Deferred weight gradients: the tail of the backward -- the audio encoder's backward, ~1200 library launches of
        # 3-10 us that leave most SMs idle -- runs AFTER the torso's; nothing depends on a weight gradient but the
        # optimiser, so the last `defer_frac` of the torso's weight-gradient work is enqueued on a side stream when the
        # torso backward ends and overlaps that tail.  B2D_DEFER_WGRAD=0 disables, a number sets the fraction.
        dw = os.environ.get("B2D_DEFER_WGRAD", "0.45")
        self.defer_frac = 0.0 if dw == "0" else float(dw)
        self._wg_total = 0.0                  # weight-gradient work of one backward (previous step), MAC count
        self._side = None
        self.seq_parallel = None              # parallel.SequenceParallelAttention, when one clip is split across GPUs
        self._step_seed = 0
        # ---- gradient layout: (param, flat offset, packed shape, unpack fn) in backward completion order
        self._grad_layout = None

    # ------------------------------------------------------------------ packed weights
    def _packed(self, key, param, fn):
        ver = (param._version, param.data_ptr())
        hit = self._wcache.get(key)
        if hit is not None and hit[0] == ver:
            return hit[1]
        ops.pack_records = recs = []
        try:
            with torch.no_grad():
                val = fn(param.detach())
        finally:
            ops.pack_records = None
        self._wcache[key] = [ver, val, param, recs]
        self._pack_table = None
        return val

    def refresh_packed(self):
        """Called at the start of every forward.  The fp16 operand copies are keyed on (`_version`, `data_ptr`) of
        their parameter, but an optimiser may update parameters without bumping `_version` (torch.optim.Adam(fused=True)
        does not), so after any forward that recorded a tape (`_weights_maybe_stale`) every copy is re-packed from its
        fp32 parameter -- in place, ONE launch of b2d_pack_conv_weights_multi for all convolutions and directions."""
        stale = self._weights_maybe_stale
        self._weights_maybe_stale = False
        entries = [(k, e) for k, e in self._wcache.items() if k != "embcat"]
        moved = [k for k, e in entries if e[0][1] != e[2].data_ptr() or any(r is None for r in e[3])]
        if moved:                                   # parameter storage replaced (.to(), assign): re-made lazily
            for k in moved:
                del self._wcache[k]
            self._pack_table = None
            entries = [(k, e) for k, e in entries if k not in set(moved)]
        if not stale and all(e[0][0] == e[2]._version for _, e in entries):
            return
        self._wcache.pop("embcat", None)
        if not entries:
            return
        if self._pack_table is None:
            recs = [r for _, e in entries for r in e[3]]
            host, blocks = ops.build_pack_table(recs, ops.L.lib().b2d_pack_block_elems())
            pinned = torch.from_numpy(host.view("uint8").reshape(-1).copy()).pin_memory()
            self._pack_table = (pinned.to(entries[0][1][2].device, non_blocking=True), len(recs), blocks, pinned)
        ops.repack_all(self._pack_table[0], self._pack_table[1], self._pack_table[2])
        for _, e in entries:
            e[0] = (e[2]._version, e[2].data_ptr())

    def w_fwd(self, conv, cin_pad=None):
        """Packed forward weights; for a Conv3d 3x3x3 (dims=3) a list of three 2-D tap groups [kt = 0, 1, 2]."""
        def fn(w):
            if w.dim() == 5:
                if w.shape[2] == 1:
                    return ops.pack_conv_weight(w[:, :, 0], cin_pad)
                return [ops.pack_conv_weight(w[:, :, kt], cin_pad) for kt in range(w.shape[2])]
            return ops.pack_conv_weight(w, cin_pad)
        return self._packed((id(conv), "f"), conv.weight, fn)

    def w_dgrad(self, conv, cin_pad=None, cout_pad=None):
        """Packed data-gradient weights; Conv3d: temporally flipped list, so that `_conv` (out[t] = sum_kt
        C(x[t + kt - 1], w[kt])) computes dx[t] = sum_kt D(dy[t - kt + 1], w[:, :, kt])."""
        def fn(w):
            if w.dim() == 5:
                if w.shape[2] == 1:
                    return ops.pack_dgrad_weight(w[:, :, 0], cin_pad, cout_pad)
                return [ops.pack_dgrad_weight(w[:, :, kt], cin_pad, cout_pad) for kt in reversed(range(w.shape[2]))]
            return ops.pack_dgrad_weight(w, cin_pad, cout_pad)
        return self._packed((id(conv), "d"), conv.weight, fn)

    # ------------------------------------------------------------------ dims=3 helpers (T frames folded into N)
    def _conv(self, x, wp, cout, k, stride, x2=None, bias=None, emb=None, residual=None, stats=False, gnb=None):
        """conv2d, or -- for a list of three temporal tap groups -- the Conv3d sum over time-shifted frames:
        the centre tap carries bias / emb / residual, the side taps accumulate in place through the residual input
        on the frame ranges that have a neighbour (zero padding in time = no contribution).
        stats: the output feeds a GroupNorm -- let the epilogue emit its pair statistics (remembered under the output's
        address for `_gn_fwd`).  gnb: ops.GnBwdFuse (the output is the dL/dy of a GroupNorm)."""
        if not isinstance(wp, list):
            st = None
            kk = k * k * (x.shape[3] + (x2.shape[3] if x2 is not None else 0))
            if (stats and self.fuse_groupnorm and kk >= self.fuse_min_k_stats
                    and ops.gn_fusable(x.shape[1] // stride, x.shape[2] // stride, cout)):
                st = self._fwd_arena.zeros((x.shape[0], cout // 2, 2), torch.float64)
            out = ops.conv2d(x, wp, cout, k, stride, x2=x2, bias=bias, emb=emb, residual=residual, stats_out=st, gnb=gnb)
            if st is not None:
                self._stats[out.data_ptr()] = (st, out)
            return out
        assert gnb is None
        T = self.T
        out = ops.conv2d(x, wp[1], cout, k, stride, x2=x2, bias=bias, emb=emb, residual=residual)
        if T > 1:
            for lo in range(0, x.shape[0], T):
                hi = lo + T
                o = out[lo:hi]
                ops.conv2d(x[lo:hi - 1], wp[0], cout, k, stride, x2=None if x2 is None else x2[lo:hi - 1],
                           residual=o[1:], out=o[1:])
                ops.conv2d(x[lo + 1:hi], wp[2], cout, k, stride, x2=None if x2 is None else x2[lo + 1:hi],
                           residual=o[:T - 1], out=o[:T - 1])
        return out

    def _wgrad(self, x, dy, gbuf, k, stride, x2=None):
        """conv2d_wgrad into a packed gradient; Conv3d gradients are packed [kt][Cout][tap][Cin]."""
        if gbuf.dim() == 2:
            return ops.conv2d_wgrad(x, dy, gbuf, k, stride, x2=x2)
        T = self.T
        ops.conv2d_wgrad(x, dy, gbuf[1], k, stride, x2=x2)
        if T > 1:
            for lo in range(0, x.shape[0], T):
                hi = lo + T
                ops.conv2d_wgrad(x[lo:hi - 1], dy[lo + 1:hi], gbuf[0], k, stride,
                                 x2=None if x2 is None else x2[lo:hi - 1])
                ops.conv2d_wgrad(x[lo + 1:hi], dy[lo:hi - 1], gbuf[2], k, stride,
                                 x2=None if x2 is None else x2[lo + 1:hi])

    def side_stream(self, device):
        if self._side is None or self._side.device != device:
            self._side = torch.cuda.Stream(device=device)
        return self._side

    def _fold(self, t):
        """(B*T, H, W, C) -> (B, T*H, W, C): GroupNorm statistics and attention run over all T*H*W positions of a sample."""
        if t is None or self.T == 1:
            return t
        n, h, w, c = t.shape
        return t.view(n // self.T, self.T * h, w, c)

    def _unfold(self, t, like):
        if t is None or self.T == 1:
            return t
        return t.view(like.shape[0], like.shape[1], like.shape[2], t.shape[3])

    def _pair_stats(self, t):
        """Pair statistics of a GroupNorm input: from the epilogue of the convolution that produced it, else one pass."""
        hit = self._stats.get(t.data_ptr())
        if hit is not None and hit[1].shape == t.shape and hit[1].stride() == t.stride() and self.T == 1:
            return hit[0]
        f = self._fold(t)
        return ops.pair_stats(f, self._fwd_arena.zeros((f.shape[0], f.shape[3] // 2, 2), torch.float64))

    def _gn_fwd(self, x, gn, x2=None, act=True, drop_p=0.0, seed=0, want_mask=False):
        """Returns (y, mean, rstd, dropout keep-bit mask or None)."""
        if not self.fuse_groupnorm:
            y, mean, rstd = ops.groupnorm_fwd(self._fold(x), gn.weight.detach(), gn.bias.detach(), x2=self._fold(x2),
                                              act=act, eps=gn.eps, drop_p=drop_p, seed=seed)
            return self._unfold(y, x), mean, rstd, None
        st1 = self._pair_stats(x)
        st2 = self._pair_stats(x2) if x2 is not None else None
        y, mean, rstd, mask = ops.groupnorm_apply(self._fold(x), st1, gn.weight.detach(), gn.bias.detach(),
                                                  x2=self._fold(x2), st2=st2, act=act, eps=gn.eps, drop_p=drop_p,
                                                  seed=seed, want_mask=want_mask)
        return self._unfold(y, x), mean, rstd, mask

    def _gnb(self, x, gn, mean, rstd, kk, x2=None, act=True, mask=None, drop_p=0.0):
        """ops.GnBwdFuse for the data-gradient convolution that produces dL/dy of act(GroupNorm([x | x2])), or None when
        the fused epilogue does not apply (dims=3 frame folding, planes the staging does not cover, dropout without
        saved mask bits, a data-gradient convolution with a reduction length kk = taps * channels below the threshold
        where the epilogue costs more than the reduce pass it replaces, A/B switch)."""
        ct = x.shape[3] + (x2.shape[3] if x2 is not None else 0)
        if (not self.fuse_groupnorm_bwd or self.T != 1 or kk < self.fuse_min_k_gnb
                or not ops.gn_fusable(x.shape[1], x.shape[2], ct)
                or (drop_p > 0.0 and mask is None)):
            return None
        ab = self._bwd_arena.zeros((x.shape[0], ct, 2), torch.float32)
        return ops.GnBwdFuse(x, x2, mean, rstd, gn.weight.detach(), gn.bias.detach(), act, ab, mask=mask,
                             keep_scale=1.0 / (1.0 - drop_p) if drop_p > 0.0 else 1.0)

    def _gn_bwd_fused(self, G, x, gn, gnb, dyh, x2=None, add_a=None, add_b=None, add_a_full=False, colsum_out=None,
                      want_colsum=False):
        """Input-gradient pass after a fused data-gradient epilogue (`gnb`) produced dyh and the channel sums."""
        if want_colsum and colsum_out is None:
            colsum_out = self._bwd_arena.zeros((x.shape[0], x.shape[-1]), torch.float32)
        dx1, dx2 = ops.groupnorm_bwd_apply(x, gn.weight.detach(), gn.bias.detach(), dyh, gnb.mean, gnb.rstd, gnb.ab,
                                           G.g(gn.weight), G.g(gn.bias), x2=x2, eps=gn.eps, add_a=add_a, add_b=add_b,
                                           add_a_full=add_a_full, colsum_out=colsum_out, overflow=self._overflow)
        if want_colsum and colsum_out is not None:
            self._colsums = {dx1.data_ptr(): (colsum_out, dx1)}
        return dx1, dx2

    def _gn_bwd(self, G, x, gn, dy, mean, rstd, x2=None, act=True, drop_p=0.0, seed=0, add_a=None, add_b=None,
                add_a_full=False, colsum_out=None, want_colsum=False):
        """want_colsum: also produce the per-sample pixel sums of dx1 (fp32 (N, C1)) inside the same kernel and remember
        them under dx1's address: the layer that consumes dx1 as a convolution's output gradient takes its bias
        gradient from there instead of re-reading dx1 (`bias_grad` in `_engine_backward`)."""
        if want_colsum and colsum_out is None and self.T == 1:
            colsum_out = torch.zeros(x.shape[0], x.shape[-1], device=x.device, dtype=torch.float32)
        dx1, dx2 = ops.groupnorm_bwd(self._fold(x), gn.weight.detach(), gn.bias.detach(), self._fold(dy), mean, rstd,
                                     G.g(gn.weight), G.g(gn.bias), x2=self._fold(x2), act=act, eps=gn.eps, drop_p=drop_p,
                                     seed=seed, add_a=self._fold(add_a), add_b=self._fold(add_b), add_a_full=add_a_full,
                                     colsum_out=colsum_out, overflow=self._overflow)
        dx1 = self._unfold(dx1, x)
        if want_colsum and colsum_out is not None:
            self._colsums = {dx1.data_ptr(): (colsum_out, dx1)}     # only the most recent one can be consumed next
        return dx1, self._unfold(dx2, x2) if x2 is not None else None

    def overflow_report(self):
        """(overflow events counted since the last backward started, back-off bits, steps skipped so far, events of the
        last finished step) -- a host read of the int32[4] device state: call it outside the step loop or one step late."""
        if self._overflow is None:
            return (0, 0.0, 0, 0)
        v = self._overflow.tolist()
        return (v[0], v[1] / 1024.0, v[2], v[3])

    def invalidate(self):
        """Forget every cached operand copy (packed 16-bit weights, concatenated embedding weights): call after
        mutating parameters behind autograd's back (`p.data.copy_()`, EMA updates, collectives on `.data`) when no
        training forward precedes the next use."""
        self._wcache.clear()
        self._pack_table = None
        self._weights_maybe_stale = False

    def emb_cat(self):
        ws = [rb.emb_layers[1].weight for rb in self.resblocks]
        bs = [rb.emb_layers[1].bias for rb in self.resblocks]
        ver = tuple((w._version, w.data_ptr()) for w in ws) + tuple((b._version, b.data_ptr()) for b in bs)
        hit = self._wcache.get("embcat")
        if hit is not None and hit[0] == ver:
            return hit[1]
        with torch.no_grad():
            val = (torch.cat([w.detach() for w in ws], 0).contiguous(), torch.cat([b.detach() for b in bs], 0).contiguous())
        self._wcache["embcat"] = (ver, val)
        return val

    def freqs(self, device):
        if self._freqs is None or self._freqs.device != device:
            self._freqs = timestep_freqs(self.model.model_channels).to(device)
        return self._freqs

    # ------------------------------------------------------------------ forward
    def forward(self, xt, cond, audio_feat, timesteps, train: bool):
        """Returns (eps_pred fp32 (N,out_ch,S,S), tape).  tape is None when train is False."""
        m = self.model
        dev = xt.device
        if xt.dim() == 5:                                 # dims=3: (B, C, T, S, S), frames folded into the batch
            B, _, T, S, _ = xt.shape
            N = B * T
        else:
            N, _, S, _ = xt.shape
            B, T = N, 1
        self.T = T
        self._stats = {}
        self._fwd_arena.begin(dev)
        self.refresh_packed()
        self._weights_maybe_stale = bool(train)
        tape = {"N": N, "S": S, "B": B, "T": T} if train else None
        drop_p = float(m.dropout) if (m.training and train) else 0.0
        self._step_seed += 1
        seed_base = (self._step_seed << 20)

        # --- conditioning vectors: time embedding MLP + batched emb_layers
        te = ops.timestep_embedding(timesteps.to(torch.int64).contiguous(), m.model_channels, self.freqs(dev))
        l0, l2 = m.time_embed[0], m.time_embed[2]
        e1 = ops.linear_fwd(te, l0.weight.detach(), l0.bias.detach())
        emb = ops.linear_fwd(e1, l2.weight.detach(), l2.bias.detach(), silu_in=True)
        wcat, bcat = self.emb_cat()
        emb_all = ops.linear_fwd(emb, wcat, bcat, silu_in=True)
        if T > 1:                                         # one row per folded frame (unet.py:262-263 broadcasts over T)
            emb_all = emb_all.repeat_interleave(T, dim=0)
        # --- input assembly (unet_audio.py:53-61) or plain NCHW -> pixel-major
        if self.is_audio:
            at = m.audio_transformer.transform[0]
            avec = ops.linear_fwd(audio_feat.contiguous(), at.weight.detach(), at.bias.detach(), relu_out=True)
            wcond = m.cond_conv_in.weight.detach().reshape(m.cond_conv_in.weight.shape[0], -1).contiguous() if m.image_cond else None
            x0 = ops.assemble_input(xt.contiguous(), cond.contiguous() if m.image_cond else None, wcond, avec, self.cin_pad)
        else:
            avec = None
            x0 = ops.nchw_to_pixel_major(xt.contiguous().view(B, xt.shape[1], T * S, S), self.cin_pad).view(N, S, S, self.cin_pad)
        if train:
            tape.update(te=te, e1=e1, emb=emb, emb_all=emb_all, avec=avec, x0=x0, cond=cond, audio_feat=audio_feat,
                        drop_p=drop_p)

        def run_layer(layer, h, skip, idx):
            seed = seed_base + idx
            if isinstance(layer, ResBlock):
                return self._res_fwd(layer, h, skip, emb_all, drop_p, seed, train)
            if isinstance(layer, AttentionBlock):
                return self._attn_fwd(layer, h, train)
            if isinstance(layer, Downsample):
                out = self._conv(h, self.w_fwd(layer.op), layer.out_channels, 3, 2, bias=layer.op.bias.detach(), stats=True)
                ctx = _Ctx("down", layer); ctx.t["x"] = h
                return out, ctx
            if isinstance(layer, Upsample):
                u = ops.upsample2x(h)
                out = self._conv(u, self.w_fwd(layer.conv), layer.out_channels, 3, 1, bias=layer.conv.bias.detach(),
                                 stats=True)
                ctx = _Ctx("up", layer); ctx.t["u"] = u
                return out, ctx
            raise RuntimeError(f"unexpected layer {type(layer)}")

        conv_in = m.input_blocks[0][0]
        h = self._conv(x0, self.w_fwd(conv_in, self.cin_pad), conv_in.out_channels, 3, 1, bias=conv_in.bias.detach(),
                       stats=True)
        hs = [h]
        idx = 0
        enc, mid, dec = [], [], []
        for block in list(m.input_blocks)[1:]:
            ctxs = []
            for layer in block:
                idx += 1
                h, ctx = run_layer(layer, h, None, idx)
                ctxs.append(ctx)
            enc.append(ctxs)
            hs.append(h)
        for layer in m.middle_block:
            idx += 1
            h, ctx = run_layer(layer, h, None, idx)
            mid.append(ctx)
        for block in m.output_blocks:
            skip = hs.pop()
            ctxs = []
            for li, layer in enumerate(block):
                idx += 1
                h, ctx = run_layer(layer, h, skip if li == 0 else None, idx)
                ctxs.append(ctx)
            dec.append(ctxs)
        if train:
            tape.update(enc=enc, mid=mid, dec=dec)
        gn, conv_out = m.out[0], m.out[2]
        hh, mean, rstd, _ = self._gn_fwd(h, gn, act=True)
        out = torch.empty(N, m.out_channels, S, S, device=dev, dtype=torch.float32)
        wo = self.w_fwd(conv_out)
        if isinstance(wo, list):        # Conv3d head: the fp32 NCHW writer cannot accumulate, so the two side taps go
            ops.conv2d(hh, wo[1], m.out_channels, 3, 1, bias=conv_out.bias.detach(), out_nchw_f32=out)   # through 3-channel
            if T > 1:                                                                                      # temporaries
                for lo in range(0, N, T):
                    hi = lo + T
                    tmp = torch.empty(T - 1, m.out_channels, S, S, device=dev, dtype=torch.float32)
                    ops.conv2d(hh[lo:hi - 1], wo[0], m.out_channels, 3, 1, out_nchw_f32=tmp)
                    out[lo + 1:hi] += tmp
                    ops.conv2d(hh[lo + 1:hi], wo[2], m.out_channels, 3, 1, out_nchw_f32=tmp)
                    out[lo:hi - 1] += tmp
        else:
            ops.conv2d(hh, wo, m.out_channels, 3, 1, bias=conv_out.bias.detach(), out_nchw_f32=out)
        if xt.dim() == 5:
            out = out.view(B, T, m.out_channels, S, S).permute(0, 2, 1, 3, 4).contiguous()
        if train:
            tape.update(head=dict(h=h, hh=hh, mean=mean, rstd=rstd))
        return out, tape

    def _res_fwd(self, rb, x, x2, emb_all, drop_p, seed, train):
        gn1, conv1 = rb.in_layers[0], rb.in_layers[2]
        gn2, conv2 = rb.out_layers[0], rb.out_layers[3]
        cout = rb.out_channels
        h1, mean1, rstd1, _ = self._gn_fwd(x, gn1, x2=x2, act=True)
        off = self.emb_off[id(rb)]
        h2 = self._conv(h1, self.w_fwd(conv1), cout, 3, 1, bias=conv1.bias.detach(), emb=emb_all[:, off:off + cout],
                        stats=True)
        h3, mean2, rstd2, mask2 = self._gn_fwd(h2, gn2, act=True, drop_p=drop_p, seed=seed, want_mask=train)
        if isinstance(rb.skip_connection, torch.nn.Identity):
            res = x
        else:
            sc = rb.skip_connection
            res = ops.conv2d(x, self.w_fwd(sc), cout, 1, 1, x2=x2, bias=sc.bias.detach())
        out = self._conv(h3, self.w_fwd(conv2), cout, 3, 1, bias=conv2.bias.detach(), residual=res, stats=True)
        ctx = None
        if train:
            ctx = _Ctx("res", rb)
            ctx.t.update(x=x, x2=x2, h1=h1, h2=h2, h3=h3, mean1=mean1, rstd1=rstd1, mean2=mean2, rstd2=rstd2,
                         drop_p=drop_p, seed=seed, mask2=mask2)
        return out, ctx

    def _attn_fwd(self, ab, x, train):
        gn = ab.norm
        C_ = ab.channels
        hn, mean, rstd, _ = self._gn_fwd(x, gn, act=False)
        qkv = ops.conv2d(hn, self.w_fwd(ab.qkv), 3 * C_, 1, 1, bias=ab.qkv.bias.detach())
        fq = self._fold(qkv)
        sp = self.seq_parallel
        rows = sp.rows(fq.shape[0], fq.shape[1] * fq.shape[2], C_ // ab.num_heads, ab.num_heads) if sp is not None else None
        a, lse = ops.attention_fwd(fq, ab.num_heads, ab.use_new_attention_order, rows=rows)   # L = T*H*W (unet.py:312)
        if rows is not None:
            sp.all_gather_rows(a, rows)                   # every rank needs the full attention output
        a = self._unfold(a, qkv)
        out = self._conv(a, self.w_fwd(ab.proj_out), C_, 1, 1, bias=ab.proj_out.bias.detach(), residual=x, stats=True)
        ctx = None
        if train:
            ctx = _Ctx("attn", ab)
            ctx.t.update(x=x, hn=hn, qkv=qkv, a=a, lse=lse, mean=mean, rstd=rstd, rows=rows)
        return out, ctx

    # ------------------------------------------------------------------ gradient storage
    def grad_layout(self):
        """[(stage, param, packed_shape)] in backward completion order; emb_layers weights are
        contiguous (resblock order) so that the batched emb GEMM's dW is one matrix."""
        if self._grad_layout is not None:
            return self._grad_layout
        m = self.model
        order = []

        def add(stage, p, shape=None):
            order.append((stage, p, tuple(shape) if shape is not None else tuple(p.shape)))

        def conv_shape(conv, cin_pad=None):
            w = conv.weight
            co, ci = w.shape[0], w.shape[1]
            taps = w.numel() // (co * ci)
            cop = max(64, _pad64(co)) if co < 64 else co
            if w.dim() == 5 and w.shape[2] > 1:           # Conv3d: [kt][Cout][kh*kw][Cin]
                return (w.shape[2], cop, (taps // w.shape[2]) * (cin_pad or ci))
            return (cop, taps * (cin_pad or ci))

        def add_layer(stage, layer):
            if isinstance(layer, ResBlock):
                for gn in (layer.in_layers[0], layer.out_layers[0]):
                    add(stage, gn.weight); add(stage, gn.bias)
                for conv in (layer.in_layers[2], layer.out_layers[3]):
                    add(stage, conv.weight, conv_shape(conv)); add(stage, conv.bias)
                if not isinstance(layer.skip_connection, torch.nn.Identity):
                    add(stage, layer.skip_connection.weight, conv_shape(layer.skip_connection))
                    add(stage, layer.skip_connection.bias)
            elif isinstance(layer, AttentionBlock):
                add(stage, layer.norm.weight); add(stage, layer.norm.bias)
                for conv in (layer.qkv, layer.proj_out):
                    add(stage, conv.weight, conv_shape(conv)); add(stage, conv.bias)
            elif isinstance(layer, Downsample):
                add(stage, layer.op.weight, conv_shape(layer.op)); add(stage, layer.op.bias)
            elif isinstance(layer, Upsample):
                add(stage, layer.conv.weight, conv_shape(layer.conv)); add(stage, layer.conv.bias)

        stage = 0
        add(stage, m.out[0].weight); add(stage, m.out[0].bias)
        add(stage, m.out[2].weight, conv_shape(m.out[2])); add(stage, m.out[2].bias, (64,))
        for block in reversed(list(m.output_blocks)):
            for layer in reversed(list(block)):
                stage += 1
                add_layer(stage, layer)
        for layer in reversed(list(m.middle_block)):
            stage += 1
            add_layer(stage, layer)
        for block in reversed(list(m.input_blocks)[1:]):
            for layer in reversed(list(block)):
                stage += 1
                add_layer(stage, layer)
        stage += 1
        conv_in = m.input_blocks[0][0]
        add(stage, conv_in.weight, conv_shape(conv_in, self.cin_pad)); add(stage, conv_in.bias)
        stage += 1
        for rb in self.resblocks:
            add(stage, rb.emb_layers[1].weight)
        for rb in self.resblocks:
            add(stage, rb.emb_layers[1].bias)
        for lin in (m.time_embed[2], m.time_embed[0]):
            add(stage, lin.weight); add(stage, lin.bias)
        if self.is_audio:
            at = m.audio_transformer.transform[0]
            add(stage, at.weight); add(stage, at.bias)
            if m.image_cond:
                add(stage, m.cond_conv_in.weight, (m.cond_conv_in.weight.shape[0], m.cond_conv_in.weight.shape[1]))
        self._grad_layout = order
        self.n_stages = stage + 1
        return order

    def torso_params(self):
        return [p for _, p, _ in self.grad_layout()]


class _Grads:
    """One flat fp32 buffer holding every torso gradient in packed layout (zeroed once per step)."""

    def __init__(self, eng: DenoiserEngine, device):
        layout = eng.grad_layout()
        self.offsets = {}
        off = 0
        self.stage_end = [0] * eng.n_stages
        for stage, p, shape in layout:
            n = 1
            for d in shape:
                n *= d
            self.offsets[id(p)] = (off, shape)
            off += (n + 3) // 4 * 4
            self.stage_end[stage] = off
        self.flat = torch.zeros(off, device=device, dtype=torch.float32)
        self.eng = eng

    def g(self, p):
        off, shape = self.offsets[id(p)]
        n = 1
        for d in shape:
            n *= d
        return self.flat[off:off + n].view(shape)

    def unpacked(self, p):
        """View of the packed gradient in the parameter's own shape (OIHW for convolutions)."""
        buf = self.g(p)
        if tuple(buf.shape) == tuple(p.shape):
            return buf
        if p.dim() == 1:                       # padded bias (head conv)
            return buf[:p.shape[0]]
        co, ci = p.shape[0], p.shape[1]
        if p.dim() == 5 and buf.dim() == 3:     # Conv3d 3x3x3: packed [kt][CoutPad][9 * CinPad] -> (O, I, kt, kh, kw)
            kt, kh, kw = p.shape[2:]
            cip = buf.shape[2] // (kh * kw)
            return buf.view(kt, buf.shape[1], kh, kw, cip)[:, :co, :, :, :ci].permute(1, 4, 0, 2, 3)
        taps = p.numel() // (co * ci)
        if p.dim() == 5:                        # Conv3d 1x1x1
            return buf.view(buf.shape[0], buf.shape[1] // taps)[:co, :ci].reshape(p.shape)
        k = int(round(taps ** 0.5)) if p.dim() == 4 else 1
        cip = buf.shape[1] // taps
        v = buf.view(buf.shape[0], taps, cip)[:co, :, :ci]         # [co][tap][ci]
        if p.dim() == 4:
            return v.reshape(co, k, k, ci).permute(0, 3, 1, 2) if (cip == ci and buf.shape[0] == co) else \
                v.contiguous().view(co, k, k, ci).permute(0, 3, 1, 2)
        if p.dim() == 3:
            return v.permute(0, 2, 1)                                # Conv1d (O, I, 1)
        return v.reshape(co, ci)


def _engine_backward(eng: DenoiserEngine, tape, dout, on_stage_done=None, defer=False):
    """Hand-scheduled backward.  dout: fp32 (N,out_ch,S,S).  Returns (_Grads, d_audio_feat or None, finish) where
    `finish` is None or a callable to run once the WHOLE backward pass has been enqueued (deferred weight gradients:
    joins the side stream, unscales their slices, then fires the stage hooks)."""
    m = eng.model
    N, B, T = tape["N"], tape["B"], tape["T"]
    eng.T = T
    dev = dout.device
    dout = dout.contiguous().float()
    if dout.dim() == 5:                                   # (B, C, T, S, S) -> frames folded into the batch
        dout = dout.permute(0, 2, 1, 3, 4).reshape(N, dout.shape[1], dout.shape[3], dout.shape[4]).contiguous()
    G = _Grads(eng, dev)
    eng._bwd_arena.begin(dev)
    if eng._overflow is None or eng._overflow.device != dev:
        eng._overflow = torch.zeros(4, device=dev, dtype=torch.int32)
    d_emb_all = eng._bwd_arena.zeros((N, eng.emb_total), torch.float32)
    stage = 0
    # fp16 gradient tensors: scale2 = [2^k, 2^-k] with max|dout| * 2^k in [64, 128); every gradient below is
    # carried times 2^k and each stage's slice of the flat fp32 buffer is unscaled as soon as the stage is done
    # (before the data-parallel reducer may all-reduce it: ranks choose different k).
    # Ranks that share one clip (sequence-parallel attention) sum partial dK / dV in 16 bits: they must use the SAME
    # scale, so the maximum is all-reduced (round-1 advisor finding: per-rank scales could differ by a factor 2 at a
    # power-of-two boundary).  The overflow state lowers the scale after a step whose gradients saturated.
    sp = eng.seq_parallel
    scale2 = ops.grad_scale(dout, state=eng._overflow, reduce_max=sp.all_reduce_max if sp is not None else None)
    inv_scale = scale2[1:]
    # ---- deferred weight gradients (see DenoiserEngine.__init__): decided per call from the work seen so far against
    # the previous step's total; only when an encoder backward follows and no gradient accumulation is pending (the
    # gradients are then handed over by the end-of-backward callback instead of through autograd, which would clone
    # the packed views -- reading them -- as soon as this function returns)
    defer_on = bool(defer) and eng.defer_frac > 0.0 and eng._wg_total > 0.0 and T == 1
    deferred, seen = [], [0.0]

    def wgrad(x, dy, gbuf, k, stride, x2=None, conv3d=True):
        cin = x.shape[3] + (x2.shape[3] if x2 is not None else 0)
        seen[0] += float(dy.numel()) * cin * k * k
        call = (eng._wgrad if conv3d else ops.conv2d_wgrad)
        if defer_on and seen[0] > (1.0 - eng.defer_frac) * eng._wg_total:
            deferred.append((call, x, dy, gbuf, k, stride, x2))
        else:
            call(x, dy, gbuf, k, stride, x2=x2)

    rr_jobs = []          # (per-sample column sums (N, C), bias gradient (C,)): summed over the samples in one launch

    def flush_rr():
        if rr_jobs:
            ops.reduce_rows_multi(rr_jobs)
            rr_jobs.clear()

    def done():
        nonlocal stage
        lo = G.stage_end[stage - 1] if stage > 0 else 0
        hi = G.stage_end[stage]
        if not defer_on:
            flush_rr()
        if hi > lo and not defer_on:                      # deferred mode: ONE unscale of the whole buffer in finish()
            ops.scale_f32_(G.flat[lo:hi], inv_scale)
        if on_stage_done is not None and not defer_on:
            on_stage_done(G, stage)
        stage += 1

    def bias_grad(dy, bias_param, cout=None, also=None):
        """bias gradient = column sums of dy over samples and pixels."""
        cout = dy.shape[3] if cout is None else cout
        hit = eng._colsums.get(dy.data_ptr())
        if hit is not None and hit[1].shape == dy.shape and hit[1].stride() == dy.stride() and hit[0].shape[1] == cout:
            tmp = hit[0]                                  # produced by the GroupNorm backward that wrote dy
        else:
            tmp = eng._bwd_arena.zeros((N, cout), torch.float32)
            ops.colsum(dy, tmp, 0, cout)
        rr_jobs.append((tmp, G.g(bias_param)))            # one batched launch per stage (or per backward), see done()
        if also is not None:
            rr_jobs.append((tmp, G.g(also)))

    # ---- head: GroupNorm -> SiLU -> conv3x3 (Cout padded to 64 for the tensor-core tiles)
    hd = tape["head"]
    gn, conv_out = m.out[0], m.out[2]
    dyp = ops.nchw_to_pixel_major(dout, 64, scale=scale2)
    wgrad(hd["hh"], dyp, G.g(conv_out.weight), 3, 1)
    bias_grad(dyp, conv_out.bias, 64)
    gnb = eng._gnb(hd["h"], gn, hd["mean"], hd["rstd"], 9 * 64, act=True)
    d_hh = eng._conv(dyp, eng.w_dgrad(conv_out, cout_pad=64), conv_out.weight.shape[1], 3, 1, gnb=gnb)
    if gnb is not None:
        dh, _ = eng._gn_bwd_fused(G, hd["h"], gn, gnb, d_hh)
    else:
        dh, _ = eng._gn_bwd(G, hd["h"], gn, d_hh, hd["mean"], hd["rstd"], act=True)
    done()

    def res_bwd(ctx, dout_, extra):
        rb, t = ctx.mod, ctx.t
        gn1, conv1 = rb.in_layers[0], rb.in_layers[2]
        gn2, conv2 = rb.out_layers[0], rb.out_layers[3]
        cout = rb.out_channels
        x, x2 = t["x"], t["x2"]
        has_skip_conv = not isinstance(rb.skip_connection, torch.nn.Identity)
        # conv2 (out_layers[3]) and the skip path share dout_
        wgrad(t["h3"], dout_, G.g(conv2.weight), 3, 1)
        bias_grad(dout_, conv2.bias, also=rb.skip_connection.bias if has_skip_conv else None)
        gnb2 = eng._gnb(t["h2"], gn2, t["mean2"], t["rstd2"], 9 * cout, act=True, mask=t.get("mask2"),
                        drop_p=t["drop_p"])
        d_h3 = eng._conv(dout_, eng.w_dgrad(conv2), cout, 3, 1, gnb=gnb2)
        off = eng.emb_off[id(rb)]
        fused_cs = eng.T == 1                              # d emb_out (unet.py:266) and conv1 bias: sums of d_h2
        if gnb2 is not None:
            d_h2, _ = eng._gn_bwd_fused(G, t["h2"], gn2, gnb2, d_h3, colsum_out=d_emb_all[:, off:off + cout])
        else:
            d_h2, _ = eng._gn_bwd(G, t["h2"], gn2, d_h3, t["mean2"], t["rstd2"], act=True, drop_p=t["drop_p"],
                                  seed=t["seed"], colsum_out=d_emb_all[:, off:off + cout] if fused_cs else None)
            if not fused_cs:
                ops.colsum(d_h2, d_emb_all[:, off:off + cout], 0, cout)
        rr_jobs.append((d_emb_all[:, off:off + cout], G.g(conv1.bias)))
        wgrad(t["h1"], d_h2, G.g(conv1.weight), 3, 1)
        cin_tot = x.shape[3] + (x2.shape[3] if x2 is not None else 0)
        gnb1 = eng._gnb(x, gn1, t["mean1"], t["rstd1"], 9 * cout, x2=x2, act=True)
        d_h1 = eng._conv(d_h2, eng.w_dgrad(conv1), cin_tot, 3, 1, gnb=gnb1)
        if has_skip_conv:
            sc = rb.skip_connection
            wgrad(x, dout_, G.g(sc.weight), 1, 1, x2=x2, conv3d=False)
            add_a = ops.conv2d(dout_, eng.w_dgrad(sc), cin_tot, 1, 1)
            full = True
        else:
            add_a, full = dout_, False
        if gnb1 is not None:
            return eng._gn_bwd_fused(G, x, gn1, gnb1, d_h1, x2=x2, add_a=add_a, add_b=extra, add_a_full=full,
                                     want_colsum=True)
        dx1, dx2 = eng._gn_bwd(G, x, gn1, d_h1, t["mean1"], t["rstd1"], x2=x2, act=True,
                               add_a=add_a, add_b=extra, add_a_full=full, want_colsum=True)
        return dx1, dx2

    def attn_bwd(ctx, dout_, extra):
        ab, t = ctx.mod, ctx.t
        C_ = ab.channels
        gn_ = ab.norm
        wgrad(t["a"], dout_, G.g(ab.proj_out.weight), 1, 1, conv3d=False)
        bias_grad(dout_, ab.proj_out.bias)
        d_a = ops.conv2d(dout_, eng.w_dgrad(ab.proj_out), C_, 1, 1)
        rows = t.get("rows")                              # sequence-parallel shard: this rank's query rows
        dqkv = ops.attention_bwd(eng._fold(t["qkv"]), eng._fold(t["a"]), eng._fold(d_a), t["lse"], ab.num_heads,
                                 ab.use_new_attention_order, rows=rows)
        if rows is not None:
            eng.seq_parallel.all_reduce_sum(dqkv)         # dQ rows from their owner, dK / dV summed over the shards
        dqkv = eng._unfold(dqkv, t["qkv"])
        wgrad(t["hn"], dqkv, G.g(ab.qkv.weight), 1, 1, conv3d=False)
        bias_grad(dqkv, ab.qkv.bias)
        gnb = eng._gnb(t["x"], gn_, t["mean"], t["rstd"], 3 * C_, act=False)
        d_hn = ops.conv2d(dqkv, eng.w_dgrad(ab.qkv), C_, 1, 1, gnb=gnb)
        if gnb is not None:
            dx, _ = eng._gn_bwd_fused(G, t["x"], gn_, gnb, d_hn, add_a=dout_, add_b=extra, want_colsum=True)
        else:
            dx, _ = eng._gn_bwd(G, t["x"], gn_, d_hn, t["mean"], t["rstd"], act=False, add_a=dout_, add_b=extra,
                                want_colsum=True)
        return dx

    def down_bwd(ctx, dout_, extra):
        layer, x = ctx.mod, ctx.t["x"]
        wgrad(x, dout_, G.g(layer.op.weight), 3, 2)
        bias_grad(dout_, layer.op.bias)
        z = ops.zero_insert2x(dout_)
        return eng._conv(z, eng.w_dgrad(layer.op), layer.channels, 3, 1, residual=extra)

    def up_bwd(ctx, dout_):
        layer, u = ctx.mod, ctx.t["u"]
        wgrad(u, dout_, G.g(layer.conv.weight), 3, 1)
        bias_grad(dout_, layer.conv.bias)
        du = eng._conv(dout_, eng.w_dgrad(layer.conv), layer.channels, 3, 1)
        return ops.upsample2x_bwd(du)

    def layer_bwd(ctx, dout_, extra):
        if ctx.kind == "res":
            return res_bwd(ctx, dout_, extra)
        if ctx.kind == "attn":
            return attn_bwd(ctx, dout_, extra), None
        if ctx.kind == "down":
            return down_bwd(ctx, dout_, extra), None
        assert extra is None
        return up_bwd(ctx, dout_), None

    # ---- decoder (reverse): collects the skip gradients for hs[0..]
    skip_grads = []
    for ctxs in reversed(tape["dec"]):
        for li in range(len(ctxs) - 1, -1, -1):
            dh, dskip = layer_bwd(ctxs[li], dh, None)
            if li == 0:
                skip_grads.append(dskip)
            done()
    # ---- middle: its first layer receives the last encoder output, which also fed the decoder
    n_enc = len(tape["enc"])
    mid = tape["mid"]
    for li in range(len(mid) - 1, -1, -1):
        dh, _ = layer_bwd(mid[li], dh, skip_grads[n_enc] if li == 0 else None)
        done()
    # ---- encoder (reverse): block i consumed hs[i] (output of block i-1 / conv_in)
    for bi in range(n_enc - 1, -1, -1):
        ctxs = tape["enc"][bi]
        for li in range(len(ctxs) - 1, -1, -1):
            dh, _ = layer_bwd(ctxs[li], dh, skip_grads[bi] if li == 0 else None)
            done()
    # ---- input convolution
    conv_in = m.input_blocks[0][0]
    x0 = tape["x0"]
    wgrad(x0, dh, G.g(conv_in.weight), 3, 1)
    bias_grad(dh, conv_in.bias)
    d_feat = None
    if eng.is_audio:
        dx0 = eng._conv(dh, eng.w_dgrad(conv_in, cin_pad=eng.cin_pad), eng.cin_pad, 3, 1)
    done()
    # ---- conditioning vectors
    wcat, _ = eng.emb_cat()
    first_w, first_b = eng.resblocks[0].emb_layers[1].weight, eng.resblocks[0].emb_layers[1].bias
    ted = wcat.shape[1]
    off_w, _ = G.offsets[id(first_w)]
    dwcat = G.flat[off_w:off_w + eng.emb_total * ted].view(eng.emb_total, ted)
    off_b, _ = G.offsets[id(first_b)]
    dbcat = G.flat[off_b:off_b + eng.emb_total]
    l0, l2 = m.time_embed[0], m.time_embed[2]
    if T > 1:                                             # emb_out was broadcast over the T frames of a sample
        d_emb_all = d_emb_all.view(B, T, -1).sum(dim=1)
    d_emb = ops.linear_bwd(tape["emb"], wcat, None, d_emb_all, dwcat, dbcat, silu_in=True)
    d_e1 = ops.linear_bwd(tape["e1"], l2.weight.detach(), None, d_emb, G.g(l2.weight), G.g(l2.bias), silu_in=True)
    ops.linear_bwd(tape["te"], l0.weight.detach(), None, d_e1, G.g(l0.weight), G.g(l0.bias), need_dx=False)
    if eng.is_audio:
        at = m.audio_transformer.transform[0]
        A = at.weight.shape[0]
        d_avec = eng._bwd_arena.zeros((N, A), torch.float32)
        cim = m.cond_conv_in.weight.shape[0] if m.image_cond else 0
        ops.assemble_input_bwd(dx0, tape["cond"].contiguous() if m.image_cond else None, m.in_channels - A - cim, cim, A,
                               G.g(m.cond_conv_in.weight) if m.image_cond else None, d_avec)
        feat = tape["audio_feat"].contiguous()
        d_feat = ops.linear_bwd(feat, at.weight.detach(), tape["avec"], d_avec, G.g(at.weight), G.g(at.bias),
                                relu_out=True)
        ops.scale_f32_(d_feat, inv_scale)
    done()
    eng._wg_total = seen[0]
    if not defer_on:
        return G, d_feat, None
    main = torch.cuda.current_stream(dev)
    side = eng.side_stream(dev)
    if deferred:
        side.wait_stream(main)
        with torch.cuda.stream(side):
            for call, x, dy, gbuf, k, stride, x2 in deferred:
                call(x, dy, gbuf, k, stride, x2=x2)

    def finish():
        # runs when autograd has enqueued everything downstream of the torso (the encoder's backward) on `main`
        with torch.cuda.stream(main):
            if deferred:
                main.wait_stream(side)
                deferred.clear()                              # the tensors may be reused by work ordered after the join
            flush_rr()
            # nothing has read the buffer yet (the stage hooks fire below, p.grad is assigned after this returns): one
            # launch unscales all of it instead of one per stage and per deferred gradient (~85 launches of 3 us)
            ops.scale_f32_(G.flat, inv_scale)
            if on_stage_done is not None:
                for st in range(eng.n_stages):
                    on_stage_done(G, st)
    return G, d_feat, finish


class _DenoiserFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, xt, cond, audio_feat, timesteps, *params):
        eng = model.engine()
        out, tape = eng.forward(xt, cond, audio_feat, timesteps, train=True)
        ctx.model, ctx.tape = model, tape
        ctx.has_audio = audio_feat is not None
        # where the gradient flows on to (the audio encoder): lets a data-parallel reducer read off the autograd graph
        # which parameters this backward will reach
        ctx.feat_grad_fn = audio_feat.grad_fn if (audio_feat is not None and audio_feat.requires_grad) else None
        return out

    @staticmethod
    def backward(ctx, dout):
        model, tape = ctx.model, ctx.tape
        if tape is None:
            raise RuntimeError("denoiser backward called on a graph built without gradients")
        eng = model.engine()
        begin = getattr(model, "_grad_begin_hook", None)
        if begin is not None:
            begin([ctx.feat_grad_fn])
        hook = getattr(model, "_grad_stage_hook", None)
        params = eng.torso_params()
        # weight gradients may be deferred past the end of this function only if something follows it (the encoder's
        # backward) and nobody reads the gradients before the end-of-backward callback: no accumulation into existing
        # .grad tensors, and the gradients are assigned there rather than returned through autograd
        defer = (ctx.feat_grad_fn is not None and all(p.grad is None for p in params)
                 and not torch.is_grad_enabled())
        G, d_feat, finish = _engine_backward(eng, tape, dout, hook, defer=defer)
        ctx.tape = None
        fin = getattr(model, "_grad_final_hook", None)
        if finish is None:
            if fin is not None:
                fin(G)
            grads = tuple(G.unpacked(p) if p.requires_grad else None for p in params)
            return (None, None, None, d_feat if ctx.has_audio else None, None) + grads
        end = getattr(model, "_grad_backward_end_hook", None)

        def at_end():
            if end is not None:
                end()                 # data-parallel: the encoder's buckets first (fixed collective order on every rank)
            finish()
            if fin is not None:
                fin(G)
            for p in params:
                if p.requires_grad:
                    p.grad = G.unpacked(p)
        torch.autograd.Variable._execution_engine.queue_callback(at_end)
        return (None, None, None, d_feat if ctx.has_audio else None, None) + (None,) * len(params)


def denoiser_apply(model, xt, cond, audio_feat, timesteps):
    if not xt.is_cuda:
        raise RuntimeError("the denoiser runs on sm_100a CUDA kernels only (no CPU fallback): move the model "
                           "and inputs to a B200 device")
    eng = model.engine()
    params = eng.torso_params()
    if torch.is_grad_enabled() and (any(p.requires_grad for p in params) or
                                    (audio_feat is not None and audio_feat.requires_grad)):
        return _DenoiserFn.apply(model, xt.float(), cond, audio_feat, timesteps, *params)
    out, _ = eng.forward(xt.float(), cond, audio_feat, timesteps, train=False)
    return out
