"""Optimiser step and densification statistics of the reference's training iteration on the B200 kernels
(csrc/optim.cu).

``FusedAdam`` takes the reference's own parameter-group list (scene/gaussian_model.py:203-217: dicts with 'params', 'lr',
'name') and does ``torch.optim.Adam(l, lr=0.0, eps=1e-15).step()`` for all groups in one kernel launch.  ``param_groups``
stays a list of dicts, so ``GaussianModel.update_learning_rate`` (:220-226) keeps working unchanged.  No CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


class FusedAdam:
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        groups = list(params)
        if groups and not isinstance(groups[0], dict):
            groups = [{"params": groups}]
        self.param_groups: List[dict] = []
        for g in groups:
            g = dict(g)
            g["params"] = list(g["params"])
            g.setdefault("lr", lr)
            self.param_groups.append(g)
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.state: Dict[torch.Tensor, dict] = {}
        self._step = 0
        n = sum(len(g["params"]) for g in self.param_groups)
        if n > 16:
            raise ValueError("FusedAdam: at most 16 parameter tensors")

    def zero_grad(self, set_to_none: bool = True):
        for g in self.param_groups:
            for p in g["params"]:
                if set_to_none:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.zero_()

    @torch.no_grad()
    def step(self):
        entries = []
        keep = []  # contiguous gradient copies must outlive the launch
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise _lib.RiggsError("FusedAdam needs contiguous fp32 CUDA parameters; riggs_b200 has no CPU path")
                st = self.state.get(p)
                if st is None:
                    st = self.state[p] = {"exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p)}
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                keep.append(grad)
                entries.append(_lib.RgAdamGroup(ptr(p), ptr(grad), ptr(st["exp_avg"]), ptr(st["exp_avg_sq"]), p.numel(),
                                                float(g["lr"])))
        if not entries:
            return
        self._step += 1
        arr = (_lib.RgAdamGroup * len(entries))(*entries)
        with torch.cuda.device(entries and self.param_groups[0]["params"][0].device):
            check(_lib.lib().rg_adam_step(len(entries), C.cast(arr, C.c_void_p), self._step, self.betas[0], self.betas[1],
                                          self.eps, stream_ptr()), "rg_adam_step")


@torch.no_grad()
def add_densification_stats(viewspace_point_tensor: torch.Tensor, radii: torch.Tensor, xyz_gradient_accum: torch.Tensor,
                            denom: torch.Tensor, max_radii2D: Optional[torch.Tensor] = None):
    """GaussianModel.add_densification_stats(viewspace_point_tensor, radii > 0) and, when max_radii2D is given, the
    max_radii2D[visibility_filter] = max(...) update of train_rig.py:335, in one launch.  ``radii`` int32 (P)."""
    g = viewspace_point_tensor.grad if viewspace_point_tensor.grad is not None else viewspace_point_tensor
    if not g.is_cuda:
        raise _lib.RiggsError("add_densification_stats needs CUDA tensors; riggs_b200 has no CPU path")
    g = g if g.is_contiguous() else g.contiguous()
    P = g.shape[0]
    if radii.dtype != torch.int32:
        radii = radii.to(torch.int32)
    for t in (xyz_gradient_accum, denom) + ((max_radii2D,) if max_radii2D is not None else ()):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P:
            raise ValueError("densification statistics must be contiguous fp32 tensors with one entry per Gaussian")
    with torch.cuda.device(g.device):
        check(_lib.lib().rg_densification_stats(P, ptr(g), ptr(radii), ptr(xyz_gradient_accum), ptr(denom),
                                                ptr(max_radii2D), stream_ptr()), "rg_densification_stats")
