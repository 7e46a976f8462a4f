"""UNetAudio -- drop-in for video-generation/diffusion/unet_audio.py:10-66.

Same constructor and forward(image, cond_image, audio, timesteps, y=None); same state_dict keys
(audio_encoder.wav2vec2.*, audio_transformer.transform.0.*, cond_conv_in.weight + the UNet torso).
The wav2vec2 encoder stays the HuggingFace module (third-party arithmetic); everything
downstream of its output -- time-mean excluded -- runs on the sm_100a kernels: Linear+ReLU projection,
the conditioning assembly (nearest resize, 1x1 cond conv, broadcast, concat) and the UNet torso.
"""
from __future__ import annotations

import collections as _collections
import os as _os

import torch
import torch.nn as nn

from .unet import UNetModel


POS_CONV_TOEPLITZ_MAX_T = 32

# Every patched piece of the third-party encoder checks the shapes its kernel is built for and hands anything else
# to the stock HF / library code.  Those hand-offs are COUNTED here per patch site (`stock_route_counts()`; bench.py
# prints them, tests assert them zero at the train.py / test.py shapes) and, with B2D_W2V_STRICT=1, raise on CUDA
# tensors instead: a step that silently ran on the library path must not pass for one that ran on these kernels.
_STOCK_ROUTES = _collections.Counter()


def stock_route_counts(reset=False):
    """{patch site: calls handed to the stock library path since import / the last reset}."""
    out = dict(_STOCK_ROUTES)
    if reset:
        _STOCK_ROUTES.clear()
    return out


def _route_to_stock(site, on_cuda):
    _STOCK_ROUTES[site] += 1
    if on_cuda and _os.environ.get("B2D_W2V_STRICT", "0") == "1":
        raise RuntimeError(f"b2d: {site}: shape / dtype not served by the sm_100a kernel and B2D_W2V_STRICT=1 "
                           "forbids the stock library path")


def _weight_normed(conv):
    """`conv.weight` of a module under torch's weight-norm parametrization, w = g * v / ||v|| with the norm over every
    dimension but `dim`, written with plain tensor ops: `torch._weight_norm`'s kernels for a norm dimension that is the
    LAST one (HF wav2vec2 normalises its positional conv over dim = 2: v (768, 48, 128), g (1, 1, 128)) walk the tensor
    with one thread per output element -- 0.35 ms forward + 0.40 ms backward per step for 4.7 M weights."""
    par = getattr(conv, "parametrizations", None)
    plist = getattr(par, "weight", None) if par is not None else None
    if plist is None or len(plist) != 1 or type(plist[0]).__name__ != "_WeightNorm":
        return conv.weight
    g, v, dim = plist.original0, plist.original1, plist[0].dim
    dim = dim if dim is None or dim >= 0 else dim + v.dim()
    if dim is None or v.dim() != 3 or g.shape != tuple(v.shape[i] if i == dim else 1 for i in range(3)):
        return conv.weight
    norm = torch.linalg.vector_norm(v.float(), dim=[i for i in range(3) if i != dim], keepdim=True)
    return v * (g / norm)


def _pos_conv_short_sequence(mod, hidden_states):
    """Wav2Vec2PositionalConvEmbedding.forward (HF transformers; third-party arithmetic, unchanged) for short
    sequences.  train.py / test.py feed 4000-sample windows = 12 frames, but the grouped Conv1d(768, 768, k=128,
    pad=64, groups=16) makes cuDNN run its direct grouped kernels over all 128 taps: 8.9 ms forward + 17 ms
    backward per step on B200 (profiles/r1_launch_list.md), almost all of it multiplying zero padding.  For
    T <= POS_CONV_TOEPLITZ_MAX_T the same sums are one batched GEMM per group against the Toeplitz matrix of the
    taps that can touch a real frame:  out[b, co, t] = sum_{ci, t'} w[co, ci, t' - t + pad] x[b, ci, t']."""
    conv = mod.conv
    B, T, C = hidden_states.shape
    K, pad, G = conv.kernel_size[0], conv.padding[0], conv.groups
    if (T > POS_CONV_TOEPLITZ_MAX_T or conv.stride[0] != 1 or conv.dilation[0] != 1 or K % 2 != 0 or pad != K // 2
            or T - 1 > pad or pad + T > K or conv.padding_mode != "zeros"):
        _route_to_stock("pos_conv", hidden_states.is_cuda)
        return mod._b2d_stock_forward(hidden_states)
    cg = C // G
    w = _weight_normed(conv)                                          # (C, C/G, K), weight-norm evaluated here
    # m[c, i, t', t] = w[c, i, t' - t + pad]: windows of the tap axis (w.unfold: U[.., j, s] = w[.., j + s]) taken at
    # j = pad - t, s = t'.  (An index gather w[:, :, idx] builds the same tensor, but its backward is an
    # index_put with accumulation: 0.7 ms per step against ~30 us for unfold's.)
    m = w.unfold(-1, T, 1)[:, :, pad - T + 1:pad + 1].flip(2).transpose(2, 3)          # (C, cg, T', T)
    m = m.view(G, cg, cg, T, T).permute(0, 2, 3, 1, 4).reshape(G, cg * T, cg * T)      # [g][(ci, t')][(co, t)]
    x = hidden_states.reshape(B, T, G, cg).permute(2, 0, 3, 1).reshape(G, B, cg * T)   # [g][b][(ci, t')]
    y = torch.bmm(x, m.to(x.dtype) if not torch.is_autocast_enabled() else m)          # (G, B, (co, t))
    y = y.view(G, B, cg, T).permute(1, 3, 0, 2).reshape(B, T, C)
    if conv.bias is not None:
        y = y + conv.bias.to(y.dtype)
    return mod.activation(y)


def _conv1d_single_channel(conv, x):
    """nn.Conv1d forward for in_channels == 1 (wav2vec2 feature_extractor.conv_layers[0]: Conv1d(1, 512, k=10,
    stride=5) over the raw waveform).  cuDNN serves this shape with its direct grouped kernels -- 8.9 ms forward and
    17 ms backward per step on B200 for 2 GFLOP (profiles/r1_launch_list.md); as a strided window view times
    the (k, C_out) weight it is one small GEMM."""
    if (conv.in_channels != 1 or conv.groups != 1 or conv.dilation[0] != 1 or conv.padding[0] != 0
            or x.dim() != 3):
        _route_to_stock("conv1d_single_channel", x.is_cuda)
        return conv._b2d_stock_forward(x)
    cols = x[:, 0].unfold(-1, conv.kernel_size[0], conv.stride[0])           # (B, T_out, k) view
    y = torch.matmul(cols, conv.weight[:, 0, :].t())                          # (B, T_out, C_out)
    if conv.bias is not None:
        y = y + conv.bias.to(y.dtype)
    return y.transpose(1, 2)


W2V_LAYER0_MAX_STEPS = 800      # b2d_w2v_layer0_*: output steps one warp keeps in registers


class _FeatureLayer0Fn(torch.autograd.Function):
    """conv(1 -> C, k = 10) -> GroupNorm(C, C) -> erf GELU of HF's `Wav2Vec2GroupNormConvLayer` as one kernel each way
    (csrc/wav2vec2_frontend.cu).  fp32 arithmetic from the fp32 waveform; output in the autocast dtype."""

    @staticmethod
    def forward(ctx, audio, w, bias, gamma, beta, stride, eps, out_dtype):
        from . import ops
        out, mean, rstd = ops.w2v_layer0_fwd(audio, w, bias, gamma, beta, stride, eps, out_dtype)
        ctx.save_for_backward(audio, w, bias, gamma, beta, mean, rstd)
        ctx.stride, ctx.eps = stride, eps
        return out

    @staticmethod
    def backward(ctx, dout):
        from . import ops
        audio, w, bias, gamma, beta, mean, rstd = ctx.saved_tensors
        dw, dbias, dgamma, dbeta = ops.w2v_layer0_bwd(audio, w, bias, gamma, beta, ctx.stride, ctx.eps,
                                                      dout.contiguous(), mean, rstd)
        return None, dw, dbias, dgamma, dbeta, None, None, None


def _feature_layer0_fused(self, hidden_states):
    """Drop-in for `Wav2Vec2GroupNormConvLayer.forward` on CUDA tensors; anything the kernel is not built for
    (other kernel width, longer windows, CPU tensors) takes the stock HF path."""
    conv, gn = self.conv, self.layer_norm
    S = hidden_states.shape[-1]
    ok = (hidden_states.is_cuda and hidden_states.dim() == 3 and hidden_states.shape[1] == 1
          and conv.kernel_size[0] == 10 and conv.padding[0] == 0 and conv.dilation[0] == 1
          and S >= 10 and 2 <= (S - 10) // conv.stride[0] + 1 <= W2V_LAYER0_MAX_STEPS and S * 4 <= 200 * 1024
          and conv.weight.dtype == torch.float32)
    if not ok:
        _route_to_stock("feature_layer0", hidden_states.is_cuda)
        return self._b2d_stock_forward(hidden_states)
    out_dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else torch.float32
    audio = hidden_states[:, 0].float().contiguous()
    return _FeatureLayer0Fn.apply(audio, conv.weight, conv.bias, gn.weight, gn.bias, conv.stride[0], gn.eps, out_dtype)


def patch_feature_layer0(wav2vec2):
    """Route the first feature-extractor layer through the fused kernels when it is the group-norm variant with an
    exact (erf) GELU -- wav2vec2-base's configuration."""
    import types
    layer0 = wav2vec2.feature_extractor.conv_layers[0]
    gn = getattr(layer0, "layer_norm", None)
    conv = getattr(layer0, "conv", None)
    if (isinstance(gn, nn.GroupNorm) and isinstance(conv, nn.Conv1d) and conv.in_channels == 1
            and gn.num_groups == gn.num_channels == conv.out_channels and gn.affine
            and getattr(wav2vec2.config, "feat_extract_activation", None) == "gelu"
            and not hasattr(layer0, "_b2d_stock_forward")):
        layer0._b2d_stock_forward = layer0.forward
        layer0.forward = types.MethodType(_feature_layer0_fused, layer0)
    return wav2vec2


SHORT_ATTENTION_MAX_TOKENS = 16      # b2d_short_attention_*: one warp per (window, head)
SHORT_ATTENTION_NAME = "b2d_short"
_short_attention_calls = [0]


class _ShortAttentionFn(torch.autograd.Function):
    """softmax(scale q k^T) -> dropout -> @ v for L <= 16, head dim 64 (csrc/short_attention.cu)."""

    @staticmethod
    def forward(ctx, q, k, v, scale, drop_p, seed):
        from . import ops
        ctx.save_for_backward(q, k, v)
        ctx.cfg = (scale, drop_p, seed)
        return ops.short_attention_fwd(q, k, v, scale, drop_p, seed)

    @staticmethod
    def backward(ctx, dout):
        from . import ops
        q, k, v = ctx.saved_tensors
        scale, drop_p, seed = ctx.cfg
        dq, dk, dv = ops.short_attention_bwd(q, k, v, dout.contiguous(), scale, drop_p, seed)
        return dq.transpose(1, 2), dk.transpose(1, 2), dv.transpose(1, 2), None, None, None


def _short_attention_interface(module, query, key, value, attention_mask, dropout=0.0, scaling=None, **kwargs):
    """HF attention-interface function (same contract as transformers' `sdpa_attention_forward`): returns
    ((B, L, H, D) output, None).  Shapes the kernel is not built for go to the library SDPA path."""
    B, H, Lq, D = query.shape
    ok = (query.is_cuda and attention_mask is None and D == 64 and Lq <= SHORT_ATTENTION_MAX_TOKENS
          and key.shape == query.shape and value.shape == query.shape and not kwargs.get("output_attentions")
          and query.dtype == key.dtype == value.dtype and query.dtype in (torch.float32, torch.bfloat16, torch.float16)
          and query.stride() == key.stride() == value.stride() and query.stride(3) == 1)
    if not ok:
        _route_to_stock("short_attention", query.is_cuda)
        from transformers.modeling_utils import ALL_ATTENTION_FUNCTIONS
        return ALL_ATTENTION_FUNCTIONS["sdpa"](module, query, key, value, attention_mask, dropout=dropout,
                                               scaling=scaling, **kwargs)
    scale = float(scaling) if scaling is not None else D ** -0.5
    _short_attention_calls[0] += 1
    seed = (torch.initial_seed() * 0x9E3779B97F4A7C15 + _short_attention_calls[0]) & 0x7FFFFFFFFFFFFFFF
    return _ShortAttentionFn.apply(query, key, value, scale, float(dropout), seed), None


def patch_short_attention(wav2vec2):
    """Select the short-sequence attention kernel for every encoder layer through transformers' attention-interface
    registry (the public hook for custom attention functions)."""
    from transformers import AttentionInterface
    AttentionInterface.register(SHORT_ATTENTION_NAME, _short_attention_interface)
    if wav2vec2.config._attn_implementation in ("sdpa", "eager"):
        wav2vec2.config._attn_implementation = SHORT_ATTENTION_NAME
    return wav2vec2


class _LayerNormFn(torch.autograd.Function):
    """nn.LayerNorm (last dimension, affine) on csrc/wav2vec2_frontend.cu's kernels: fp32 arithmetic and fp32 output,
    which is what the stock module gives for an fp32 input and, under autocast, for a 16-bit one."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        from . import ops
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        y, mean, rstd = ops.layernorm_fwd(x2, weight, bias, eps)
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.x_dtype = x.dtype
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        from . import ops
        x2, weight, mean, rstd = ctx.saved_tensors
        dx, dg, db = ops.layernorm_bwd(x2, weight, mean, rstd, dy.reshape(x2.shape).float().contiguous())
        return dx.view(dy.shape).to(ctx.x_dtype), dg, db, None


def _layer_norm_fused(self, x):
    """Drop-in for `nn.LayerNorm.forward` on CUDA tensors; anything else takes the stock path."""
    ok = (x.is_cuda and x.dim() >= 2 and len(self.normalized_shape) == 1 and self.elementwise_affine
          and self.weight is not None and self.bias is not None and self.weight.dtype == torch.float32
          and x.shape[-1] // 32 in (2, 4, 8, 16, 24, 32) and x.shape[-1] % 32 == 0 and x.numel() > 0
          and (x.dtype == torch.float32 or (x.dtype in (torch.bfloat16, torch.float16) and torch.is_autocast_enabled("cuda"))))
    if not ok:
        _route_to_stock("layer_norm", x.is_cuda)
        return self._b2d_stock_forward(x)
    return _LayerNormFn.apply(x, self.weight, self.bias, self.eps)


def patch_layer_norms(wav2vec2):
    """Route every affine nn.LayerNorm of the encoder through the fused kernels (row N4)."""
    import types
    for mod in wav2vec2.modules():
        if type(mod) is nn.LayerNorm and not hasattr(mod, "_b2d_stock_forward"):
            mod._b2d_stock_forward = mod.forward
            mod.forward = types.MethodType(_layer_norm_fused, mod)
    return wav2vec2


class _LinearFn(torch.autograd.Function):
    """nn.Linear as the stock module computes it (under autocast: in the autocast dtype, one library GEMM with the bias
    in its epilogue), with the bias gradient from `b2d_colsum_rows16` in fp32 instead of torch's 16-bit `sum` + cast and
    dx / dW taken from the gradient GEMMs' fp32 accumulators instead of their 16-bit roundings + casts."""

    @staticmethod
    def forward(ctx, x, weight, bias, w16=None, b16=None):
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        xb = x.to(dt)
        # w16 / b16: this step's 16-bit copies from the encoder's cast cache (one multi-tensor copy per forward instead
        # of two cast kernels per Linear); they are refreshed in place by the next forward, so they are kept as plain
        # attributes -- autograd's version check is for tensors whose VALUES may have changed, these hold the same values
        # until the optimiser has stepped
        wb = weight.to(dt) if w16 is None else w16
        with torch.autocast("cuda", enabled=False):
            y = torch.nn.functional.linear(xb, wb, bias.to(dt) if b16 is None else b16)
        ctx.save_for_backward(xb)
        ctx.wb = wb
        ctx.dts = (x.dtype, weight.dtype, bias.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        from . import ops
        (xb,), wb = ctx.saved_tensors, ctx.wb
        dx = dw = db = None
        dy2 = dy.reshape(-1, dy.shape[-1])
        if dy2.dtype != xb.dtype:
            dy2 = dy2.to(xb.dtype)
        dy2 = dy2.contiguous()
        with torch.autocast("cuda", enabled=False):
            # fp32 consumers (the fp32 residual stream, the fp32 parameters) get the GEMM's fp32 accumulator directly:
            # the stock path rounds it to 16 bits and casts it back with one more kernel per gradient
            def mm(a, b, want):
                if want == torch.float32 and a.dtype != torch.float32:
                    return torch.mm(a, b, out_dtype=torch.float32)
                return torch.mm(a, b).to(want)
            if ctx.needs_input_grad[0]:
                dx = mm(dy2, wb, ctx.dts[0]).view(xb.shape)
            if ctx.needs_input_grad[1]:
                dw = mm(dy2.t(), xb.reshape(-1, xb.shape[-1]), ctx.dts[1])
        if ctx.needs_input_grad[2]:
            db = ops.colsum_rows16(dy2).to(ctx.dts[2])
        return dx, dw, db, None, None


_CAST_CACHE_ACTIVE = [False]      # True while Wav2Vec2Encoder.forward runs with a refreshed cast cache


def _linear_fused(self, x):
    """Drop-in for `nn.Linear.forward` in training on CUDA with a 16-bit compute dtype; anything else: stock path."""
    dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    ok = (x.is_cuda and self.bias is not None and torch.is_grad_enabled() and self.weight.requires_grad
          and dt in (torch.bfloat16, torch.float16) and x.dtype in (torch.float32, torch.bfloat16, torch.float16)
          and self.out_features % 2 == 0 and x.numel() > 0)
    if not ok:
        # eval / no-grad calls (the sampler) and fp32 runs are the stock module by design: the patch only replaces
        # the 16-bit training backward.  Counted separately so that a training step can assert on "linear_train".
        _route_to_stock("linear_train" if (x.is_cuda and torch.is_grad_enabled() and self.weight.requires_grad
                                           and dt in (torch.bfloat16, torch.float16)) else "linear_eval_or_fp32", False)
        return self._b2d_stock_forward(x)
    w16 = getattr(self, "_b2d_w16", None) if _CAST_CACHE_ACTIVE[0] else None
    if w16 is not None and w16.dtype == dt and w16.device == x.device:
        return _LinearFn.apply(x, self.weight, self.bias, w16, self._b2d_b16)
    return _LinearFn.apply(x, self.weight, self.bias)


def patch_linears(wav2vec2):
    """Route the encoder's nn.Linear layers through `_LinearFn` (bias gradient on this library's kernel; row N4)."""
    import types
    for mod in wav2vec2.modules():
        if type(mod) is nn.Linear and not hasattr(mod, "_b2d_stock_forward"):
            mod._b2d_stock_forward = mod.forward
            mod.forward = types.MethodType(_linear_fused, mod)
    return wav2vec2


def _mask_hidden_states_no_sync(self, hidden_states, mask_time_indices=None, attention_mask=None):
    """HF `Wav2Vec2Model._mask_hidden_states` (SpecAugment) without its four host<->device synchronisations per train
    step: the stock code uploads the numpy mask from pageable memory (`torch.tensor(mask, device=...)` -- a stream
    sync) and applies it with boolean-mask indexing (`nonzero` -- a device->host read, again in the backward), which
    drains the launch queue twice per step and leaves the GPU idle while the host re-fills it.  Same mask (the same
    `_compute_mask_indices` calls, so the numpy RNG stream is consumed identically), uploaded through pinned memory
    and applied with `torch.where`: identical values, nothing waits."""
    from transformers.models.wav2vec2 import modeling_wav2vec2 as W
    if not getattr(self.config, "apply_spec_augment", True):
        return hidden_states
    if not hidden_states.is_cuda:
        _route_to_stock("spec_augment", False)
        return self._b2d_stock_mask_hidden_states(hidden_states, mask_time_indices=mask_time_indices,
                                                  attention_mask=attention_mask)
    batch_size, sequence_length, hidden_size = hidden_states.size()

    def upload(mask_np):
        host = torch.from_numpy(mask_np.astype("bool", copy=False)).pin_memory()
        return host.to(hidden_states.device, non_blocking=True)

    if mask_time_indices is not None:
        m = mask_time_indices.to(torch.bool)
        hidden_states = torch.where(m[:, :, None], self.masked_spec_embed.to(hidden_states.dtype), hidden_states)
    elif self.config.mask_time_prob > 0 and self.training:
        m = upload(W._compute_mask_indices((batch_size, sequence_length), mask_prob=self.config.mask_time_prob,
                                           mask_length=self.config.mask_time_length, attention_mask=attention_mask,
                                           min_masks=self.config.mask_time_min_masks))
        hidden_states = torch.where(m[:, :, None], self.masked_spec_embed.to(hidden_states.dtype), hidden_states)
    if self.config.mask_feature_prob > 0 and self.training:
        f = upload(W._compute_mask_indices((batch_size, hidden_size), mask_prob=self.config.mask_feature_prob,
                                           mask_length=self.config.mask_feature_length,
                                           min_masks=self.config.mask_feature_min_masks))
        hidden_states = torch.where(f[:, None, :], torch.zeros((), dtype=hidden_states.dtype, device=hidden_states.device),
                                    hidden_states)
    return hidden_states


def patch_spec_augment(wav2vec2):
    """Replace the SpecAugment masking step of a HF Wav2Vec2Model by the synchronisation-free restatement above."""
    import types
    if not hasattr(wav2vec2, "_b2d_stock_mask_hidden_states"):
        wav2vec2._b2d_stock_mask_hidden_states = wav2vec2._mask_hidden_states
        wav2vec2._mask_hidden_states = types.MethodType(_mask_hidden_states_no_sync, wav2vec2)


class _ToChannelsLastFn(torch.autograd.Function):
    """(B, C, T) contiguous -> the (B, C, 1, T) channels-last view of its (B, T, C) transpose, and back for the gradient,
    on the tiled 16-bit transpose kernel (torch's strided copy: ~1 TB/s)."""

    @staticmethod
    def forward(ctx, h):
        from . import ops
        B, C, T = h.shape
        return ops.transpose16(h.contiguous()).view(B, 1, T, C).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        from . import ops
        B, C, _, T = dy.shape
        d = dy.permute(0, 2, 3, 1).contiguous().view(B, T, C)          # a view when dy is channels-last, as it is here
        return ops.transpose16(d)


def _feature_encoder_channels_last(self, input_values):
    """Drop-in for `Wav2Vec2FeatureEncoder.forward` (third-party; row N4).  The stock layers 1..6 are nn.Conv1d on
    (B, C, T) tensors: torch hands cuDNN contiguous NCW tensors, cuDNN's 16-bit kernels are NHWC, and every convolution
    call is wrapped in `nchwToNhwc` / `nhwcToNchw` conversions -- 1.6 ms of conversions around 1.2 ms of convolution per
    train64 step, forward + backward (profiles/r2_launch_list.md).  Here the activation is converted ONCE after layer 0
    to a (B, C, 1, T) channels-last tensor (physically (B, T, C)) and the same library convolution is called as conv2d
    with a 1 x k kernel: cuDNN runs its NHWC kernels on it directly, forward and backward, and the (B, C, T) view
    returned at the end is the contiguous (B, T, C) tensor HF transposes it to."""
    layers = list(self.conv_layers)
    ok = (input_values.is_cuda and len(layers) > 1 and not getattr(self, "gradient_checkpointing", False)
          and all(type(l).__name__ == "Wav2Vec2NoLayerNormConvLayer" and isinstance(l.conv, nn.Conv1d)
                  and l.conv.padding[0] == 0 and l.conv.dilation[0] == 1 and l.conv.groups == 1
                  and l.conv.padding_mode == "zeros" for l in layers[1:]))
    if not ok:
        _route_to_stock("feature_convs", input_values.is_cuda)
        return self._b2d_stock_forward(input_values)
    h = layers[0](input_values[:, None])
    if h.dtype in (torch.bfloat16, torch.float16):
        h = _ToChannelsLastFn.apply(h)
    else:
        h = h.unsqueeze(2).contiguous(memory_format=torch.channels_last)
    for layer in layers[1:]:
        conv = layer.conv
        h = nn.functional.conv2d(h, conv.weight.unsqueeze(2), conv.bias, stride=(1, conv.stride[0]))
        h = layer.activation(h)
    return h.squeeze(2)


def patch_feature_convs(wav2vec2):
    import types
    fe = wav2vec2.feature_extractor
    if not hasattr(fe, "_b2d_stock_forward"):
        fe._b2d_stock_forward = fe.forward
        fe.forward = types.MethodType(_feature_encoder_channels_last, fe)
    return wav2vec2


def patch_pos_conv(wav2vec2):
    """Route the positional conv embedding of a HF Wav2Vec2Model through _pos_conv_short_sequence and the
    single-channel waveform conv through _conv1d_single_channel (same arithmetic, library GEMMs)."""
    import types
    pce = wav2vec2.encoder.pos_conv_embed
    if not hasattr(pce, "_b2d_stock_forward"):
        pce._b2d_stock_forward = pce.forward
        pce.forward = types.MethodType(_pos_conv_short_sequence, pce)
    conv0 = wav2vec2.feature_extractor.conv_layers[0].conv
    if conv0.in_channels == 1 and not hasattr(conv0, "_b2d_stock_forward"):
        conv0._b2d_stock_forward = conv0.forward
        conv0.forward = types.MethodType(_conv1d_single_channel, conv0)
    return wav2vec2


class Wav2Vec2Encoder(nn.Module):
    """unet_audio.py:10-18.  `from_pretrained` needs the checkpoint in the HF cache; pass
    `pretrained=False` (or set B2D_WAV2VEC2_RANDOM_INIT=1) for a random-init wav2vec2-base."""

    def __init__(self, model_name="facebook/wav2vec2-base-960h",
                 cache_dir="/proj/vondrick/aa4870/hf-model-checkpoints", pretrained=None):
        super().__init__()
        import os
        from transformers import Wav2Vec2Config, Wav2Vec2Model
        if pretrained is None:
            pretrained = os.environ.get("B2D_WAV2VEC2_RANDOM_INIT", "0") != "1"
        if pretrained:
            self.wav2vec2 = Wav2Vec2Model.from_pretrained(model_name, cache_dir=cache_dir)
        else:
            self.wav2vec2 = Wav2Vec2Model(Wav2Vec2Config())
        if os.environ.get("B2D_WAV2VEC2_STOCK", "0") != "1":     # A/B switch: leave the HF module untouched
            patch_pos_conv(self.wav2vec2)
            patch_feature_layer0(self.wav2vec2)
            patch_short_attention(self.wav2vec2)
            patch_spec_augment(self.wav2vec2)
            if os.environ.get("B2D_W2V_LAYERNORM", "1") != "0":
                patch_layer_norms(self.wav2vec2)
            if os.environ.get("B2D_W2V_LINEAR", "1") != "0":
                patch_linears(self.wav2vec2)
            if os.environ.get("B2D_W2V_CHANNELS_LAST", "1") != "0":
                patch_feature_convs(self.wav2vec2)

        self._cast_cache = None
        self._cast_cache_on = os.environ.get("B2D_W2V_CAST_CACHE", "1") != "0"

    def _refresh_cast_cache(self, dt, device):
        """16-bit copies of the patched Linear layers' weights and biases for this forward: views of one flat buffer,
        refreshed by ONE multi-tensor copy (autocast casts each fp32 weight and bias with a kernel of its own on every
        call: ~150 launches of 3-8 us per train step)."""
        mods = [m for m in self.wav2vec2.modules()
                if type(m) is nn.Linear and hasattr(m, "_b2d_stock_forward") and m.bias is not None]
        key = (dt, device, tuple(id(m.weight) for m in mods))
        if self._cast_cache is None or self._cast_cache[0] != key:
            if not mods or any(m.weight.device != device or m.weight.dtype != torch.float32 for m in mods):
                self._cast_cache = (key, None, None)
            else:
                src = [t for m in mods for t in (m.weight, m.bias)]
                sizes = [(t.numel() + 7) // 8 * 8 for t in src]                  # 16-byte aligned views
                flat = torch.empty(sum(sizes), dtype=dt, device=device)
                dst, off = [], 0
                for t, n in zip(src, sizes):
                    dst.append(flat[off:off + t.numel()].view(t.shape))
                    off += n
                for i, m in enumerate(mods):
                    m._b2d_w16, m._b2d_b16 = dst[2 * i], dst[2 * i + 1]
                self._cast_cache = (key, src, dst)
        _, src, dst = self._cast_cache
        if src is None:
            return False
        with torch.no_grad():
            torch._foreach_copy_(dst, src)
        return True

    def forward(self, audio_input):
        x = audio_input.get("input_values") if hasattr(audio_input, "get") else None      # dict or HF BatchFeature
        use = (self._cast_cache_on and self.training and torch.is_grad_enabled() and x is not None and x.is_cuda
               and torch.is_autocast_enabled("cuda") and hasattr(torch, "_foreach_copy_"))
        if use:
            use = self._refresh_cast_cache(torch.get_autocast_dtype("cuda"), x.device)
        prev = _CAST_CACHE_ACTIVE[0]
        _CAST_CACHE_ACTIVE[0] = bool(use)
        try:
            return self.wav2vec2(**audio_input, output_hidden_states=True).last_hidden_state
        finally:
            _CAST_CACHE_ACTIVE[0] = prev


class AudioFeatureTransformer(nn.Module):
    """unet_audio.py:21-30 (parameter holder; the Linear+ReLU runs in the engine)."""

    def __init__(self, input_dim, output_dim):
        super().__init__()
        self.transform = nn.Sequential(nn.Linear(input_dim, output_dim), nn.ReLU())


class UNetAudio(UNetModel):
    def __init__(self, image_size, in_channels, model_channels, out_channels, num_res_blocks,
                 attention_resolutions, image_cond=True, im_cond_input_ch=3, im_cond_output_ch=64, dropout=0.1,
                 channel_mult=(1, 2, 4), conv_resample=True, dims=2, num_classes=None, use_checkpoint=False,
                 use_fp16=False, num_heads=1, num_head_channels=-1, num_heads_upsample=-1,
                 use_scale_shift_norm=False, resblock_updown=False, use_new_attention_order=False,
                 audio_feature_dim=512, projected_audio_dim=256, audio_encoder=None):
        super().__init__(image_size, in_channels + projected_audio_dim + (im_cond_output_ch if image_cond else 0),
                         model_channels, out_channels, num_res_blocks, attention_resolutions, dropout, channel_mult,
                         conv_resample, dims, num_classes, use_checkpoint, use_fp16, num_heads, num_head_channels,
                         num_heads_upsample, use_scale_shift_norm, resblock_updown, use_new_attention_order)
        self.audio_encoder = audio_encoder if audio_encoder is not None else Wav2Vec2Encoder()
        self.audio_transformer = AudioFeatureTransformer(audio_feature_dim, projected_audio_dim)
        self.projected_audio_dim = projected_audio_dim
        self.image_size = image_size
        self.image_cond = image_cond
        if self.image_cond:
            self.cond_conv_in = nn.Conv2d(in_channels=im_cond_input_ch, out_channels=im_cond_output_ch,
                                          kernel_size=1, bias=False)
        self._cached_audio = None

    def encode_audio(self, audio):
        """wav2vec2 features averaged over time: (N, audio_feature_dim).  Loop-invariant in sampling
        (test.py:57-65 recomputes it every step); `sample_images` hoists it."""
        return self.audio_encoder(audio).mean(dim=1)

    def forward(self, image, cond_image, audio, timesteps, y=None, audio_features=None):
        if image.shape[-1] != self.image_size or image.shape[-2] != self.image_size:
            # the reference's .expand(-1,-1,image_size,image_size) + cat fails for other sizes
            raise RuntimeError(f"Sizes of tensors must match: image is {tuple(image.shape[-2:])}, "
                               f"model image_size is {self.image_size}")
        feat = audio_features if audio_features is not None else self.encode_audio(audio)
        return self._torso(image, cond_image if self.image_cond else None, feat.float(), timesteps)
