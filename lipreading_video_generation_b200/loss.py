"""MSELoss -- drop-in for the `nn.MSELoss()` criterion of train.py:103,132 on an sm_100a reduction kernel
(warp shuffle -> block -> deterministic last-block pass) with a fused elementwise backward."""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib as L

_ws: dict = {}


def _workspace(device):
    key = (str(device), torch.cuda.current_stream().cuda_stream)
    if key not in _ws:
        nbytes = L.lib().b2d_mse_workspace_bytes(0)
        _ws[key] = torch.zeros(nbytes, device=device, dtype=torch.uint8)   # counter starts at zero, self-resets
    return _ws[key]


class _MSEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        L.require_cuda(pred, target)
        if pred.shape != target.shape:
            raise RuntimeError(f"MSELoss: shape mismatch {tuple(pred.shape)} vs {tuple(target.shape)}")
        p, t = pred.contiguous().float(), target.contiguous().float()
        loss = torch.empty((), device=pred.device, dtype=torch.float32)
        ws = _workspace(pred.device)
        L.check(L.lib().b2d_mse_fwd(p.data_ptr(), t.data_ptr(), p.numel(), ws.data_ptr(), ws.numel(),
                                    loss.data_ptr(), L.stream_ptr()), "b2d_mse_fwd")
        ctx.save_for_backward(p, t)
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        p, t = ctx.saved_tensors
        g = grad_out.contiguous().float()
        dp = torch.empty_like(p)
        L.check(L.lib().b2d_mse_bwd(p.data_ptr(), t.data_ptr(), g.data_ptr(), p.numel(), dp.data_ptr(),
                                    L.stream_ptr()), "b2d_mse_bwd")
        return (dp if ctx.needs_input_grad[0] else None), (-dp if ctx.needs_input_grad[1] else None)


def mse_loss(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    return _MSEFn.apply(pred, target)


class MSELoss(nn.Module):
    """`criterion = MSELoss(); loss = criterion(noise_pred, noise)` (train.py:103,132), reduction='mean'."""

    def __init__(self, size_average=None, reduce=None, reduction: str = "mean"):
        super().__init__()
        if reduction != "mean" or size_average is not None or reduce is not None:
            raise NotImplementedError("only the default reduction='mean' of train.py:103 is provided")
        self.reduction = reduction

    def forward(self, input, target):
        return mse_loss(input, target)
