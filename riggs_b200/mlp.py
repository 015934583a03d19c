"""The two per-Gaussian MLPs of the skinning path (WeightMLP, DeformMLP = detail_net) on the tensor cores.

``FusedSkipMLP`` wraps one of the torch modules of ``skeleton_warp`` (whose parameters stay the source of truth: same
names as the reference's, so checkpoints load and torch optimisers update them) and evaluates it with the library's
tcgen05 kernels (csrc/mlp.cu) instead of ~20 eager cuBLAS / element-wise launches: one kernel carries 128 points through
all nine layers with the activations resident in tensor memory.  bf16 operands, fp32 accumulation -- see DESIGN.md for
the tolerance against the reference's fp32 evaluation.

There is no torch fallback inside this class: without the CUDA library it raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import check, f32c, ptr, stream_ptr


class FusedSkipMLP:
    """module: skeleton_warp.WeightMLP (sigmoid head, parameters ``linear.{i}``, ``weight_predict``) or
    skeleton_warp.DeformMLP (linear head, ``linear.{i}``, ``gaussian_warp``; its second input -- the pose, identical
    for every point -- is folded into the biases of layers 0 and 5)."""

    def __init__(self, module: torch.nn.Module):
        self.module = module
        self.layers = list(module.linear)
        if hasattr(module, "weight_predict"):
            self.head, self.out_act = module.weight_predict, 1
        else:
            self.head, self.out_act = module.gaussian_warp, 0
        self.multires = len(module.embed.freqs)
        self.emb_cols = 3 * (1 + 2 * self.multires)
        self.out_ch = self.head.out_features
        self.extra = self.layers[0].in_features - self.emb_cols  # pose columns (0 for WeightMLP)
        dev = self.head.weight.device
        nbytes = int(_lib.lib().rg_mlp_packed_bytes(self.out_ch))
        self.packed = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        self._packed_versions = None
        self.repack()

    # ---- weights ---------------------------------------------------------------------------------------------
    def _weights(self):
        return [l.weight for l in self.layers] + [self.head.weight]

    def repack(self, force: bool = False) -> None:
        """bf16 chunk images of the current weights; cheap no-op while no weight tensor has been modified in place."""
        ws = self._weights()
        versions = tuple((w.data_ptr(), w._version) for w in ws)
        if not force and versions == self._packed_versions:
            return
        wc = [f32c(w.detach()) for w in ws]
        n = len(wc)
        ptrs = (C.c_void_p * n)(*[w.data_ptr() for w in wc])
        lds = (C.c_int * n)(*[w.shape[1] for w in wc])
        emb0 = (C.c_int * n)(*[0] * n)
        act0 = [0] * n
        act0[5] = self.emb_cols + self.extra  # layer 5 sees [encoding, pose, activations]
        act0 = (C.c_int * n)(*act0)
        with torch.cuda.device(self.packed.device):
            check(_lib.lib().rg_mlp_pack(self.emb_cols, self.out_ch, ptrs, lds, emb0, act0, self.packed.data_ptr(), stream_ptr()),
                  "rg_mlp_pack")
        self._keep = wc
        self._packed_versions = versions

    def biases(self, pose: Optional[torch.Tensor] = None) -> torch.Tensor:
        """(9, 256) fp32: the layers' biases with the pose's contribution folded into layers 0 and 5."""
        dev = self.packed.device
        b = torch.zeros(9, 256, device=dev)
        for i, l in enumerate(self.layers):
            b[i] = l.bias.detach()
        b[8, :self.out_ch] = self.head.bias.detach()
        if self.extra:
            if pose is None:
                raise _lib.RiggsError("this MLP has a per-frame input (the pose): pass it")
            p = pose.detach().reshape(-1).float()
            lo, hi = self.emb_cols, self.emb_cols + self.extra
            b[0] += self.layers[0].weight.detach()[:, lo:hi] @ p
            b[5] += self.layers[5].weight.detach()[:, lo:hi] @ p
        return b

    # ---- evaluation ------------------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, pose: Optional[torch.Tensor] = None, save: bool = False):
        """x (N,3) -> (N,out_ch).  save=True also returns the state the backward needs."""
        if not x.is_cuda:
            raise _lib.RiggsError("FusedSkipMLP needs CUDA tensors; riggs_b200 has no CPU path")
        self.repack()
        x = f32c(x.detach())
        N = x.shape[0]
        out = torch.empty(N, self.out_ch, device=x.device)
        bias = self.biases(pose)
        acts = emb = None
        n_pad = 0
        if save:
            n_pad = (N + 63) // 64 * 64
            acts = torch.empty(8, 256, n_pad, dtype=torch.bfloat16, device=x.device)
            emb = torch.zeros(64, n_pad, dtype=torch.bfloat16, device=x.device)
        with torch.cuda.device(x.device):
            check(_lib.lib().rg_mlp_forward(N, self.multires, self.out_ch, self.out_act, ptr(x), self.packed.data_ptr(),
                                            bias.data_ptr(), ptr(out), ptr(acts), ptr(emb), n_pad, stream_ptr()), "rg_mlp_forward")
        if save:
            return out, (x, acts, emb, n_pad, bias)
        return out

    __call__ = forward
