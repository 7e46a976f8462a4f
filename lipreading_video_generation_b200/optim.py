"""Adam -- drop-in for `optimizer = Adam(model.parameters(), 1e-2)` / `optimizer.step()` (train.py:102,134) on one
sm_100a multi-tensor kernel: every parameter tensor of the model (344 tensors, 110.7 M fp32
values at the train.py config) is updated by a single launch of `b2d_adam_step`, 28 bytes of HBM traffic per
parameter.  State keys (`step`, `exp_avg`, `exp_avg_sq`) and `state_dict()` layout are torch.optim.Adam's, so
optimiser checkpoints move between the two.  No CPU fallback: CPU parameters raise."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L

CHUNK = 65536          # elements per CTA
_CHUNK_DTYPE = np.dtype([("param", "<u8"), ("grad", "<u8"), ("exp_avg", "<u8"), ("exp_avg_sq", "<u8"), ("n", "<i8")])


def build_chunk_table(ptrs, chunk: int = CHUNK) -> np.ndarray:
    """Host logic of the multi-tensor launch: split each (param_ptr, grad_ptr, m_ptr, v_ptr, numel) into runs of
    <= `chunk` consecutive fp32 elements; returns a structured array laid out as `B2dAdamChunk[]` (include/b2d.h)."""
    a = np.asarray(ptrs, dtype=np.int64).reshape(-1, 5)
    n = a[:, 4]
    per = (n + chunk - 1) // chunk                        # chunks per tensor (0 for an empty tensor)
    which = np.repeat(np.arange(len(a)), per)             # tensor index of every chunk
    first = np.repeat(np.cumsum(per) - per, per)
    lo = (np.arange(int(per.sum())) - first) * chunk      # first element of the chunk inside its tensor
    t = np.empty(len(which), dtype=_CHUNK_DTYPE)
    for j, name in enumerate(("param", "grad", "exp_avg", "exp_avg_sq")):
        t[name] = (a[which, j] + 4 * lo).astype(np.uint64)
    t["n"] = np.minimum(chunk, n[which] - lo)
    return t


class Adam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (amsgrad / maximize / capturable / differentiable not provided: train.py:102 uses
    none of them).  Gradients that are not fp32-contiguous are made so first (one copy); parameters must be fp32
    contiguous CUDA tensors."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *,
                 maximize=False, foreach=None, capturable=False, differentiable=False, fused=None, skip_on_overflow=None):
        if amsgrad or maximize or capturable or differentiable:
            raise NotImplementedError("Adam: amsgrad / maximize / capturable / differentiable are not used by "
                                      "train.py:102 and are not provided")
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0) or weight_decay < 0.0:
            raise ValueError("Adam: invalid hyper-parameter")          # torch/optim/adam.py raises the same way
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False,
                                      maximize=False, foreach=None, capturable=False, differentiable=False,
                                      fused=None))
        self._tables = {}          # group index -> (key, device table, pinned host table)
        # skip_on_overflow: a model whose engine counts saturated 16-bit gradients on the device (or the int32 tensor
        # itself): the update becomes a no-op for a step whose backward overflowed -- the reference's low-precision
        # recipe (fp16_util.py:222-233), decided on the device without a host synchronisation
        self._overflow_src = skip_on_overflow

    def _table(self, gi, ptrs, device):
        key = tuple(ptrs)
        ent = self._tables.get(gi)
        if ent is not None and ent[0] == key:
            return ent[1]
        host = build_chunk_table(ptrs)
        pinned = torch.from_numpy(host.view(np.uint8).reshape(-1).copy()).pin_memory()
        dev = pinned.to(device, non_blocking=True)
        # the previous pinned buffer may still be the source of an in-flight copy: keep it until the next rebuild
        self._tables[gi] = (key, dev, pinned, ent[2] if ent is not None else None)
        return dev

    def _overflow_ptr(self, device):
        src = self._overflow_src
        if src is None:
            return None
        t = src if isinstance(src, torch.Tensor) else getattr(src.engine(), "_overflow", None)
        return t.data_ptr() if (t is not None and t.device == device) else None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = L.lib()
        for gi, group in enumerate(self.param_groups):
            beta1, beta2 = group["betas"]
            by_step = {}
            touched = []
            steps = []
            for p in group["params"]:
                g = p.grad
                if g is None:
                    continue
                L.require_cuda(p, g)
                if g.is_sparse:
                    raise RuntimeError("Adam does not support sparse gradients")
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("Adam: parameters must be contiguous fp32 tensors (the state_dict layout)")
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.to(torch.float32).contiguous()
                    p.grad = g
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                if st["step"].is_cuda:              # state loaded from torch's fused / capturable Adam: keep it on the host
                    st["step"] = st["step"].detach().to("cpu", torch.float32)
                steps.append(st["step"])
                k = int(st["step"]) + 1
                by_step.setdefault((k, p.device), []).append(
                    (p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel()))
                touched.append(p)
            for si, ((k, device), ptrs) in enumerate(sorted(by_step.items(), key=lambda kv: kv[0][0])):
                with torch.cuda.device(device):
                    table = self._table((gi, si), ptrs, device)
                    n_chunks = table.numel() // _CHUNK_DTYPE.itemsize
                    L.check(lib.b2d_adam_step_guarded(table.data_ptr(), n_chunks, float(group["lr"]), float(beta1),
                                                      float(beta2), float(group["eps"]), float(group["weight_decay"]),
                                                      1.0 - beta1 ** k, 1.0 - beta2 ** k, self._overflow_ptr(device),
                                                      L.stream_ptr()), "b2d_adam_step_guarded")
            if touched:
                torch._foreach_add_(steps, 1)
                torch.autograd.graph.increment_version(touched)     # packed fp16 operand copies are keyed on _version
        return loss
