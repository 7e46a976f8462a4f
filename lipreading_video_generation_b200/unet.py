"""UNetModel -- drop-in for video-generation/diffusion/unet.py:408-675 on sm_100a kernels.

The module tree below exists to own the parameters under the reference's exact
state_dict keys (time_embed.{0,2}, input_blocks.N.M.in_layers.{0,2}, ... out.{0,2});
the arithmetic is done by `DenoiserEngine` (engine.py), a hand-scheduled forward /
backward over the C-ABI kernels, exposed to autograd as ONE torch.autograd.Function.
Constructor signature, forward signature and error behaviour follow the reference.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn


class GroupNorm32(nn.GroupNorm):
    """Parameter holder for utils.py:54-56 (the fp32-statistics GroupNorm)."""


def zero_module(module: nn.Module) -> nn.Module:
    """utils.py:105-111"""
    for p in module.parameters():
        p.detach().zero_()
    return module


def normalization(channels: int) -> GroupNorm32:
    """utils.py:130-137"""
    return GroupNorm32(32, channels)


def conv_nd(dims: int, *args, **kwargs) -> nn.Module:
    """utils.py:59-69 (parameter holder: Conv2d / Conv3d own the OIHW / OIDHW tensors under the reference keys)."""
    if dims == 2:
        return nn.Conv2d(*args, **kwargs)
    if dims == 3:
        return nn.Conv3d(*args, **kwargs)
    raise ValueError(f"unsupported dimensions: {dims}")


class TimestepEmbedSequential(nn.Sequential):
    """unet.py:78-90 (container only; dispatch happens in the engine)."""


class Upsample(nn.Module):
    """unet.py:93-122"""

    def __init__(self, channels, use_conv, dims=2, out_channels=None):
        super().__init__()
        if not use_conv:
            raise NotImplementedError("conv_resample=False is not on the train.py/test.py path")
        self.channels, self.out_channels, self.use_conv, self.dims = channels, out_channels or channels, use_conv, dims
        self.conv = conv_nd(dims, self.channels, self.out_channels, 3, padding=1)


class Downsample(nn.Module):
    """unet.py:125-152"""

    def __init__(self, channels, use_conv, dims=2, out_channels=None):
        super().__init__()
        if not use_conv:
            raise NotImplementedError("conv_resample=False is not on the train.py/test.py path")
        self.channels, self.out_channels, self.use_conv, self.dims = channels, out_channels or channels, use_conv, dims
        self.op = conv_nd(dims, self.channels, self.out_channels, 3, stride=2 if dims != 3 else (1, 2, 2), padding=1)  # unet.py:141


class ResBlock(nn.Module):
    """unet.py:155-268"""

    def __init__(self, channels, emb_channels, dropout, out_channels=None, use_conv=False,
                 use_scale_shift_norm=False, dims=2, use_checkpoint=False, up=False, down=False):
        super().__init__()
        if use_scale_shift_norm or up or down or use_conv:
            raise NotImplementedError("use_scale_shift_norm / resblock_updown / use_conv are not on the train.py path")
        self.channels, self.emb_channels, self.dropout = channels, emb_channels, dropout
        self.out_channels = out_channels or channels
        self.use_checkpoint = use_checkpoint
        self.in_layers = nn.Sequential(normalization(channels), nn.SiLU(),
                                       conv_nd(dims, channels, self.out_channels, 3, padding=1))
        self.emb_layers = nn.Sequential(nn.SiLU(), nn.Linear(emb_channels, self.out_channels))
        self.out_layers = nn.Sequential(normalization(self.out_channels), nn.SiLU(), nn.Dropout(p=dropout),
                                        zero_module(conv_nd(dims, self.out_channels, self.out_channels, 3, padding=1)))
        if self.out_channels == channels:
            self.skip_connection = nn.Identity()
        else:
            self.skip_connection = conv_nd(dims, channels, self.out_channels, 1)


class AttentionBlock(nn.Module):
    """unet.py:271-317"""

    def __init__(self, channels, num_heads=1, num_head_channels=-1, use_checkpoint=False,
                 use_new_attention_order=False):
        super().__init__()
        self.channels = channels
        if num_head_channels == -1:
            self.num_heads = num_heads
        else:
            assert channels % num_head_channels == 0, (
                f"q,k,v channels {channels} is not divisible by num_head_channels {num_head_channels}")
            self.num_heads = channels // num_head_channels
        self.use_checkpoint = use_checkpoint
        self.use_new_attention_order = use_new_attention_order
        self.norm = normalization(channels)
        self.qkv = nn.Conv1d(channels, channels * 3, 1)
        self.proj_out = zero_module(nn.Conv1d(channels, channels, 1))


class UNetModel(nn.Module):
    """unet.py:408-675.  Same constructor; forward(x, timesteps, y=None) -> (N, out_channels, S, S) fp32."""

    def __init__(self, image_size, in_channels, model_channels, out_channels, num_res_blocks,
                 attention_resolutions, dropout=0, channel_mult=(1, 2, 4, 8), conv_resample=True, dims=2,
                 num_classes=None, use_checkpoint=False, use_fp16=False, num_heads=1, num_head_channels=-1,
                 num_heads_upsample=-1, use_scale_shift_norm=False, resblock_updown=False,
                 use_new_attention_order=False):
        super().__init__()
        if dims not in (2, 3):
            raise NotImplementedError("dims=1 is never used by the reference's callers")
        if num_classes is not None or use_fp16 or resblock_updown or use_scale_shift_norm or not conv_resample:
            raise NotImplementedError("num_classes / use_fp16 / resblock_updown / use_scale_shift_norm / "
                                      "conv_resample=False are never used by train.py / test.py")
        if model_channels % 64 != 0:
            raise NotImplementedError("model_channels must be a multiple of 64 for the tcgen05 tiles")
        if num_heads_upsample == -1:
            num_heads_upsample = num_heads
        self.image_size, self.in_channels, self.model_channels = image_size, in_channels, model_channels
        self.dims = dims
        self.out_channels, self.num_res_blocks = out_channels, num_res_blocks
        self.attention_resolutions, self.dropout, self.channel_mult = attention_resolutions, dropout, channel_mult
        self.conv_resample, self.num_classes, self.use_checkpoint = conv_resample, num_classes, use_checkpoint
        self.dtype = torch.float32
        self.num_heads, self.num_head_channels, self.num_heads_upsample = num_heads, num_head_channels, num_heads_upsample

        time_embed_dim = model_channels * 4
        self.time_embed = nn.Sequential(nn.Linear(model_channels, time_embed_dim), nn.SiLU(),
                                        nn.Linear(time_embed_dim, time_embed_dim))
        ch = input_ch = int(channel_mult[0] * model_channels)
        self.input_blocks = nn.ModuleList([TimestepEmbedSequential(conv_nd(dims, in_channels, ch, 3, padding=1))])
        input_block_chans = [ch]
        ds = 1
        for level, mult in enumerate(channel_mult):
            for _ in range(num_res_blocks):
                layers = [ResBlock(ch, time_embed_dim, dropout, out_channels=int(mult * model_channels), dims=dims,
                                   use_checkpoint=use_checkpoint)]
                ch = int(mult * model_channels)
                if ds in attention_resolutions:
                    layers.append(AttentionBlock(ch, use_checkpoint=use_checkpoint, num_heads=num_heads,
                                                 num_head_channels=num_head_channels,
                                                 use_new_attention_order=use_new_attention_order))
                self.input_blocks.append(TimestepEmbedSequential(*layers))
                input_block_chans.append(ch)
            if level != len(channel_mult) - 1:
                self.input_blocks.append(TimestepEmbedSequential(Downsample(ch, conv_resample, dims=dims, out_channels=ch)))
                input_block_chans.append(ch)
                ds *= 2
        self.middle_block = TimestepEmbedSequential(
            ResBlock(ch, time_embed_dim, dropout, dims=dims, use_checkpoint=use_checkpoint),
            AttentionBlock(ch, use_checkpoint=use_checkpoint, num_heads=num_heads,
                           num_head_channels=num_head_channels, use_new_attention_order=use_new_attention_order),
            ResBlock(ch, time_embed_dim, dropout, dims=dims, use_checkpoint=use_checkpoint))
        self.output_blocks = nn.ModuleList([])
        for level, mult in list(enumerate(channel_mult))[::-1]:
            for i in range(num_res_blocks + 1):
                ich = input_block_chans.pop()
                layers = [ResBlock(ch + ich, time_embed_dim, dropout, out_channels=int(model_channels * mult), dims=dims,
                                   use_checkpoint=use_checkpoint)]
                ch = int(model_channels * mult)
                if ds in attention_resolutions:
                    layers.append(AttentionBlock(ch, use_checkpoint=use_checkpoint, num_heads=num_heads_upsample,
                                                 num_head_channels=num_head_channels,
                                                 use_new_attention_order=use_new_attention_order))
                if level and i == num_res_blocks:
                    layers.append(Upsample(ch, conv_resample, dims=dims, out_channels=ch))
                    ds //= 2
                self.output_blocks.append(TimestepEmbedSequential(*layers))
        self.out = nn.Sequential(normalization(ch), nn.SiLU(),
                                 zero_module(conv_nd(dims, input_ch, out_channels, 3, padding=1)))
        self._engine = None

    # -- engine plumbing -------------------------------------------------
    def engine(self):
        from .engine import DenoiserEngine
        if self._engine is None or self._engine.model is not self:
            self._engine = DenoiserEngine(self)
        return self._engine

    def __deepcopy__(self, memo):
        """`copy.deepcopy(model)` (an EMA copy): the engine -- operand caches and per-module tables keyed on THIS model's
        modules -- is not copied; the copy builds its own on first use."""
        import copy
        eng, self._engine = self._engine, None
        try:
            new = self.__class__.__new__(self.__class__)
            memo[id(self)] = new
            new.__dict__ = copy.deepcopy(self.__dict__, memo)
        finally:
            self._engine = eng
        return new

    def _torso(self, xt, cond, audio_feat, timesteps):
        from .engine import denoiser_apply
        return denoiser_apply(self, xt, cond, audio_feat, timesteps)

    def forward(self, x, timesteps, y=None):
        """unet.py:646-675.  x: (N, in_channels, S, S) fp32 on an sm_100a device; (B, in_channels, T, S, S) under
        dims=3 (Conv3d 3x3x3 = three time-shifted 2-D tap groups on the frame-folded batch, GroupNorm and attention
        over T*S*S tokens per sample, Downsample stride (1,2,2), Upsample (T, 2H, 2W): engine.py)."""
        if x.dim() != self.dims + 2:
            raise RuntimeError(f"expected a {self.dims + 2}-D input for dims={self.dims}, got {tuple(x.shape)}")
        assert (y is not None) == (self.num_classes is not None), \
            "must specify y if and only if the model is class-conditional"
        return self._torso(x, None, None, timesteps)

    def convert_to_fp16(self):
        raise NotImplementedError("use_fp16 path is never called by train.py/test.py; compute is bf16/fp32-accumulate")

    def convert_to_fp32(self):
        return None


def timestep_freqs(dim: int, max_period: int = 10000) -> torch.Tensor:
    """The frequency table of utils.py:150-152, built with the reference's own torch ops (bit-exact)."""
    half = dim // 2
    return torch.exp(-math.log(max_period) * torch.arange(start=0, end=half, dtype=torch.float32) / half)
