"""ORACLE helper (test infrastructure): import the REAL reference modules from /root/reference.

Works only where /root/reference exists (the build container); never used at run time on the GPU box.
Follows SURVEY.md section 8(c): Wav2Vec2Model.from_pretrained is patched to a random-init
Wav2Vec2Model(Wav2Vec2Config()) (no network), train.py / test.py are never imported (import-time
nvidia-smi, absolute-path pickles, decord)."""
from __future__ import annotations

import os
import sys

REF_DIR = "/root/reference/video-generation/diffusion"


def available() -> bool:
    return os.path.isdir(REF_DIR)


def load():
    """Returns a namespace with the reference modules: unet, unet_audio, linear_noise_scheduler, noise_scheduler, utils."""
    if not available():
        raise RuntimeError(f"{REF_DIR} not present")
    sys.dont_write_bytecode = True
    import transformers
    from transformers import Wav2Vec2Config, Wav2Vec2Model
    if not getattr(Wav2Vec2Model, "_b2d_patched", False):
        Wav2Vec2Model.from_pretrained = classmethod(lambda cls, *a, **k: cls(Wav2Vec2Config()))
        Wav2Vec2Model._b2d_patched = True
    shadow = {}
    names = ["unet", "unet_audio", "linear_noise_scheduler", "noise_scheduler", "utils", "fp16_util"]
    for n in names:                       # the reference uses flat imports ("from unet import ...")
        shadow[n] = sys.modules.pop(n, None)
    sys.path.insert(0, REF_DIR)
    try:
        import importlib
        mods = {n: importlib.import_module(n) for n in names}
    finally:
        sys.path.remove(REF_DIR)
        for n in names:
            sys.modules.pop(n, None)
            if shadow[n] is not None:
                sys.modules[n] = shadow[n]

    class NS:
        pass
    ns = NS()
    for n, mod in mods.items():
        setattr(ns, n, mod)
    return ns
