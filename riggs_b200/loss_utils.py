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


# ---- L1 + SSIM (train_rig.py:508-509; utils/loss_utils.py:36-77) ---------------------------------------------------
_SSIM_SCRATCH: Dict[tuple, torch.Tensor] = {}


def _ssim_call(network_output: torch.Tensor, gt: torch.Tensor, lambda_dssim: float):
    """-> (loss3 = [loss, l1, ssim] device tensor, grad of the loss w.r.t. network_output)."""
    if not network_output.is_cuda:
        raise _lib.RiggsError("the image loss needs CUDA tensors; riggs_b200 has no CPU path")
    if network_output.shape != gt.shape or network_output.dim() not in (3, 4):
        raise ValueError(f"image loss: expected two (C,H,W) or (B,C,H,W) tensors of one shape, got "
                         f"{tuple(network_output.shape)} and {tuple(gt.shape)}")
    a, b = f32c(network_output.detach()), f32c(gt.detach())
    H, W = a.shape[-2], a.shape[-1]
    C = a.numel() // (H * W)  # batch folds into channels: every (H, W) plane is filtered on its own
    L = _lib.lib()
    nbytes = int(L.rg_ssim_loss_scratch_bytes(C, H, W))
    key = (str(a.device), torch.cuda.current_stream(a.device).cuda_stream, nbytes)
    sc = _SSIM_SCRATCH.get(key)
    if sc is None:
        sc = _SSIM_SCRATCH[key] = torch.empty(nbytes, dtype=torch.uint8, device=a.device)
    loss3 = torch.empty(3, dtype=torch.float32, device=a.device)
    grad = torch.empty_like(a)
    with torch.cuda.device(a.device):
        check(L.rg_ssim_l1_loss(C, H, W, ptr(a), ptr(b), float(lambda_dssim), ptr(loss3), ptr(grad), ptr(sc), sc.numel(),
                                stream_ptr()), "rg_ssim_l1_loss")
    return loss3, grad


def image_loss_and_grad(network_output: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2):
    """(loss, d loss / d network_output, l1, ssim) of the trainer's image loss
    ``(1 - lambda_dssim) * l1_loss + lambda_dssim * (1 - ssim)``, outside autograd (two kernel launches)."""
    with torch.no_grad():
        loss3, grad = _ssim_call(network_output, gt, lambda_dssim)
    return loss3[0], grad, loss3[1], loss3[2]


class _ImageLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, network_output, gt, lambda_dssim, which):
        loss3, grad = _ssim_call(network_output, gt, lambda_dssim)
        ctx.save_for_backward(grad)
        ctx.sign = -1.0 if which == 2 else 1.0  # which = 2 returns ssim itself: d ssim = -(d loss at lambda = 1)
        return loss3[which].clone()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * (g * ctx.sign), None, None, None


def image_loss(network_output: torch.Tensor, gt: torch.Tensor, lambda_dssim: float = 0.2) -> torch.Tensor:
    """``(1.0 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1.0 - ssim(image, gt))`` as one differentiable op."""
    return _ImageLoss.apply(network_output, gt, float(lambda_dssim), 0)


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11, size_average: bool = True) -> torch.Tensor:
    """utils/loss_utils.py ``ssim`` (differentiable in img1).  Only the reference's defaults are implemented."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("riggs_b200.loss_utils.ssim implements window_size=11, size_average=True")
    return _ImageLoss.apply(img1, img2, 1.0, 2)
