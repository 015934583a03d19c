"""Image losses of the reference's utils/loss_utils.py that sit between the rasterizer's forward and backward.

Only ``l1_loss`` is provided: with autograd it is seven image-sized element-wise launches per iteration in the
reference; here the loss and its gradient come out of one kernel (csrc/loss.cu).  There is no CPU path.
"""
from __future__ import annotations

from typing import Dict

import torch

from . import _lib
from ._lib import check, f32c, ptr, stream_ptr

_SCRATCH: Dict[tuple, torch.Tensor] = {}


def _scratch(dev: torch.device) -> torch.Tensor:
    key = (str(dev), torch.cuda.current_stream(dev).cuda_stream)
    t = _SCRATCH.get(key)
    if t is None:
        t = torch.zeros(int(_lib.lib().rg_l1_loss_scratch_bytes()), dtype=torch.uint8, device=dev)
        _SCRATCH[key] = t
    return t


class _L1Loss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, network_output, gt):
        if not network_output.is_cuda:
            raise _lib.RiggsError("l1_loss needs CUDA tensors; riggs_b200 has no CPU path")
        if network_output.shape != gt.shape:
            raise ValueError(f"l1_loss: shapes differ: {tuple(network_output.shape)} vs {tuple(gt.shape)}")
        a, b = f32c(network_output), f32c(gt)
        loss = torch.empty((), dtype=torch.float32, device=a.device)
        grad = torch.empty_like(a)
        sc = _scratch(a.device)
        with torch.cuda.device(a.device):
            check(_lib.lib().rg_l1_loss(a.numel(), ptr(a), ptr(b), ptr(loss), ptr(grad), ptr(sc), sc.numel(), stream_ptr()),
                  "rg_l1_loss")
        ctx.save_for_backward(grad)
        ctx.need_gt = gt.requires_grad
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        d = grad * g
        return d, (-d if ctx.need_gt else None)


def l1_loss_and_grad(network_output: torch.Tensor, gt: torch.Tensor):
    """(mean |network_output - gt|, d loss / d network_output) from one kernel launch, outside autograd: a training
    step whose only loss is L1 can feed the gradient straight to ``network_output.backward(grad)`` and skip the
    image-sized multiply by the upstream 1.0 that autograd would run."""
    if not network_output.is_cuda:
        raise _lib.RiggsError("l1_loss needs CUDA tensors; riggs_b200 has no CPU path")
    if network_output.shape != gt.shape:
        raise ValueError(f"l1_loss: shapes differ: {tuple(network_output.shape)} vs {tuple(gt.shape)}")
    with torch.no_grad():
        a, b = f32c(network_output.detach()), f32c(gt.detach())
        loss = torch.empty((), dtype=torch.float32, device=a.device)
        grad = torch.empty_like(a)
        sc = _scratch(a.device)
        with torch.cuda.device(a.device):
            check(_lib.lib().rg_l1_loss(a.numel(), ptr(a), ptr(b), ptr(loss), ptr(grad), ptr(sc), sc.numel(), stream_ptr()),
                  "rg_l1_loss")
    return loss, grad


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """mean |network_output - gt| (utils/loss_utils.py l1_loss); the gradient is produced by the same kernel."""
    return _L1Loss.apply(network_output, gt)
