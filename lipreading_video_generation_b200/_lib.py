"""ctypes binding of libb200diff.so (the C ABI declared in include/b2d.h).

There is deliberately no fallback: if the shared library is missing or a call
returns non-zero, a RuntimeError is raised (north_star: "no CPU fallback").
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200diff.so")

_lib = None

c_void_p, c_int, c_int64, c_float, c_size_t = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t


class ConvArgs(C.Structure):
    _fields_ = [
        ("x", c_void_p), ("ldx", c_int64), ("C1", c_int),
        ("x2", c_void_p), ("ldx2", c_int64), ("C2", c_int),
        ("N", c_int), ("H", c_int), ("W", c_int),
        ("ksize", c_int), ("stride", c_int),
        ("w", c_void_p), ("Cout", c_int), ("CoutPad", c_int),
        ("bias", c_void_p), ("emb", c_void_p), ("ldemb", c_int64),
        ("residual", c_void_p), ("ldres", c_int64),
        ("out", c_void_p), ("ldo", c_int64),
        ("out_nchw_f32", c_void_p),
        ("gn_stats_out", c_void_p),
        ("gnb_x1", c_void_p), ("gnb_ld1", c_int64), ("gnb_C1", c_int),
        ("gnb_x2", c_void_p), ("gnb_ld2", c_int64),
        ("gnb_mean", c_void_p), ("gnb_rstd", c_void_p), ("gnb_gamma", c_void_p), ("gnb_beta", c_void_p),
        ("gnb_act", c_int), ("gnb_mask", c_void_p), ("gnb_keep_scale", c_float),
        ("gnb_ab", c_void_p),
    ]


class WgradArgs(C.Structure):
    _fields_ = [
        ("x", c_void_p), ("ldx", c_int64), ("C1", c_int),
        ("x2", c_void_p), ("ldx2", c_int64), ("C2", c_int),
        ("N", c_int), ("H", c_int), ("W", c_int),
        ("ksize", c_int), ("stride", c_int),
        ("dy", c_void_p), ("lddy", c_int64), ("Cout", c_int),
        ("dw", c_void_p),
    ]


class GnArgs(C.Structure):
    _fields_ = [
        ("x1", c_void_p), ("ld1", c_int64), ("C1", c_int),
        ("x2", c_void_p), ("ld2", c_int64), ("C2", c_int),
        ("N", c_int), ("HW", c_int),
        ("gamma", c_void_p), ("beta", c_void_p),
        ("eps", c_float), ("act", c_int), ("drop_p", c_float), ("seed", C.c_uint64),
        ("overflow", c_void_p),
    ]


# name -> (restype, argtypes); every symbol include/b2d.h declares must appear here
SIGNATURES = {
    "b2d_version": (c_int, []),
    "b2d_last_error": (C.c_char_p, []),
    "b2d_check_device": (c_int, []),
    "b2d_launch_count": (C.c_longlong, []),
    "b2d_q_sample": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int64, c_int64,
                             c_void_p, c_void_p]),
    "b2d_ddpm_step": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_float, c_float, c_float, c_float, c_int,
                              c_void_p, c_void_p, c_int64, c_void_p]),
    "b2d_ddpm_v2_step": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_float, c_float, c_float, c_void_p,
                                 c_void_p, c_int64, c_void_p]),
    "b2d_cosine_step": (c_int, [c_void_p, c_void_p, c_void_p, c_float, c_float, c_float, c_int, c_void_p, c_void_p,
                                c_int64, c_void_p]),
    "b2d_ddim_step": (c_int, [c_void_p, c_void_p, c_float, c_float, c_float, c_float, c_void_p, c_void_p, c_int64,
                              c_void_p]),
    "b2d_mse_workspace_bytes": (c_size_t, [c_int64]),
    "b2d_mse_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_size_t, c_void_p, c_void_p]),
    "b2d_mse_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "b2d_conv2d": (c_int, [C.POINTER(ConvArgs), c_void_p]),
    "b2d_conv2d_gn_fusable": (c_int, [c_int, c_int, c_int]),
    "b2d_conv2d_wgrad": (c_int, [C.POINTER(WgradArgs), c_void_p]),
    "b2d_pair_stats": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "b2d_groupnorm_apply": (c_int, [C.POINTER(GnArgs), c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    "b2d_groupnorm_bwd_apply": (c_int, [C.POINTER(GnArgs), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                        c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_int64,
                                        c_void_p, c_void_p, c_void_p]),
    "b2d_attention_fwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p,
                                  c_void_p]),
    "b2d_pack_conv_weight": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int,
                                     c_void_p]),
    "b2d_attention_bwd_ws_floats": (c_int64, [c_int, c_int, c_int, c_int]),
    "b2d_attention_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int, c_int,
                                  c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "b2d_groupnorm_workspace_bytes": (c_size_t, [c_int, c_int]),
    "b2d_groupnorm_fwd": (c_int, [C.POINTER(GnArgs), c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "b2d_groupnorm_fwd_static": (c_int, [C.POINTER(GnArgs), c_void_p, c_void_p, c_void_p, c_void_p]),
    "b2d_groupnorm_bwd": (c_int, [C.POINTER(GnArgs), c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
                                  c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_void_p,
                                  c_void_p, c_size_t, c_void_p]),
    "b2d_assemble_input": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 9 + [c_void_p, c_void_p]),
    "b2d_assemble_input_bwd": (c_int, [c_void_p, c_void_p] + [c_int] * 9 + [c_void_p, c_void_p, c_void_p]),
    "b2d_upsample2x": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "b2d_upsample2x_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "b2d_zero_insert2x": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "b2d_nchw_to_pixel_major": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "b2d_grad_scale": (c_int, [c_void_p, c_int64, c_float, c_void_p, c_void_p]),
    "b2d_scale_f32": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "b2d_absmax": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "b2d_grad_scale_finalize": (c_int, [c_void_p, c_float, c_void_p, c_void_p]),
    "b2d_w2v_layer0_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                   c_float, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "b2d_w2v_layer0_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                   c_float, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p]),
    "b2d_short_attention_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int,
                                        c_int, c_float, c_float, C.c_uint64, c_void_p, c_void_p]),
    "b2d_short_attention_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int, c_int,
                                        c_int, c_int, c_int, c_float, c_float, C.c_uint64, c_void_p, c_void_p, c_void_p,
                                        c_void_p]),
    "b2d_colsum": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "b2d_reduce_rows": (c_int, [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "b2d_layernorm_fwd": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b2d_layernorm_bwd": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p]),
    "b2d_colsum_rows16": (c_int, [c_void_p, c_int, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "b2d_transpose16": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "b2d_attention_fwd_splitkv_ws_floats": (c_int64, [c_int, c_int, c_int, c_int, c_int]),
    "b2d_attention_fwd_splitkv": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int64,
                                          c_void_p, c_int64, c_void_p, c_void_p]),
    "b2d_reduce_rows_multi": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p]),
    "b2d_timestep_embedding": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "b2d_linear_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b2d_linear_bwd": (c_int, [c_void_p] * 7 + [c_int] * 5 + [c_void_p]),
    "b2d_pack_block_elems": (c_int, []),
    "b2d_pack_conv_weights_multi": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "b2d_audio_window_max_samples": (c_int, []),
    "b2d_audio_window": (c_int, [c_void_p] * 10 + [c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b2d_frame_gather": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "b2d_attention_fwd_rows": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int64,
                                       c_void_p, c_void_p]),
    "b2d_attention_bwd_rows": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int, c_int,
                                       c_int, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "b2d_adam_step": (c_int, [c_void_p, c_int, C.c_double, C.c_double, C.c_double, c_float, c_float, C.c_double, C.c_double,
                              c_void_p]),
    "b2d_adam_step_guarded": (c_int, [c_void_p, c_int, C.c_double, C.c_double, C.c_double, c_float, c_float, C.c_double,
                                      C.c_double, c_void_p, c_void_p]),
}


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raise loudly when it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m lipreading_video_generation_b200.build` "
                "(there is no CPU / PyTorch fallback for the sm_100a kernels)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().b2d_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("libb200diff kernels need CUDA tensors on an sm_100a device (no CPU fallback)")
