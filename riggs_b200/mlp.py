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
        emb0, act0 = self._col_args()
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
    def _col_args(self):
        n = 9
        act0 = [0] * n
        act0[5] = self.emb_cols + self.extra  # layer 5 sees [encoding, pose, activations]
        return (C.c_int * n)(*[0] * n), (C.c_int * n)(*act0)

    def forward(self, x: torch.Tensor, pose: Optional[torch.Tensor] = None, save: bool = False):
        """x (N,3) -> (N,out_ch), no autograd.  save=True also returns the state ``backward`` needs."""
        if not x.is_cuda:
            raise _lib.RiggsError("FusedSkipMLP needs CUDA tensors; riggs_b200 has no CPU path")
        # under CUDA-graph capture the packing is recorded too: a replay then packs whatever the weights hold by then
        self.repack(force=torch.cuda.is_current_stream_capturing())
        x = f32c(x.detach())
        N = x.shape[0]
        out = torch.empty(N, self.out_ch, device=x.device)
        bias = self.biases(pose)
        acts = emb = masks = None
        n_pad = 0
        if save:
            n_pad = (N + 127) // 128 * 128
            acts = torch.empty(8, 256, n_pad, dtype=torch.bfloat16, device=x.device)
            emb = torch.empty(64, n_pad, dtype=torch.bfloat16, device=x.device)
            emb[self.emb_cols:].zero_()  # the kernel writes the rows of the encoding only
            masks = torch.empty(8, 8, n_pad, dtype=torch.int32, device=x.device)
        with torch.cuda.device(x.device):
            check(_lib.lib().rg_mlp_forward(N, self.multires, self.out_ch, self.out_act, ptr(x), self.packed.data_ptr(),
                                            bias.data_ptr(), ptr(out), ptr(acts), ptr(emb), ptr(masks), n_pad, stream_ptr()),
                  "rg_mlp_forward")
        if save:
            return out, (N, acts, emb, masks, n_pad, pose.detach().reshape(-1).float() if pose is not None else None)
        return out

    def backward(self, g_out: torch.Tensor, out: torch.Tensor, state):
        """dL/d out (N,out_ch) -> gradients of the module's parameters, in the order of ``parameters()``:
        [linear.0.weight, linear.0.bias, ..., head.weight, head.bias]."""
        N, acts, emb, masks, n_pad, pose = state
        dev = g_out.device
        ws = self._weights()
        dws = [torch.zeros_like(w, dtype=torch.float32) for w in ws]
        db = torch.zeros(9, 256, device=dev)
        dz = torch.empty(9, 256, n_pad, dtype=torch.bfloat16, device=dev)
        n = len(ws)
        ptrs = (C.c_void_p * n)(*[d.data_ptr() for d in dws])
        lds = (C.c_int * n)(*[w.shape[1] for w in ws])
        emb0, act0 = self._col_args()
        g = f32c(g_out)
        with torch.cuda.device(dev):
            check(_lib.lib().rg_mlp_backward(N, self.out_ch, self.out_act, self.emb_cols, ptr(g), ptr(f32c(out)) if self.out_act else None,
                                             self.packed.data_ptr(), ptr(acts), ptr(emb), ptr(masks), ptr(dz), n_pad, ptrs, lds, emb0, act0,
                                             db.data_ptr(), stream_ptr()), "rg_mlp_backward")
        if self.extra:  # the pose columns saw the same input for every point: their gradient is d_bias (x) pose
            lo, hi = self.emb_cols, self.emb_cols + self.extra
            dws[0][:, lo:hi] = torch.outer(db[0], pose)
            dws[5][:, lo:hi] = torch.outer(db[5], pose)
        grads = []
        for i in range(8):
            grads += [dws[i], db[i].clone()]
        grads += [dws[8], db[8, :self.out_ch].clone()]
        return grads

    def parameters(self):
        ps = []
        for l in self.layers:
            ps += [l.weight, l.bias]
        return ps + [self.head.weight, self.head.bias]

    def __call__(self, x: torch.Tensor, pose: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Differentiable w.r.t. the module's parameters when autograd is recording (x and the pose are inputs the
        reference detaches, skeleton_warp.py:131, 152); plain inference otherwise."""
        if not torch.is_grad_enabled() or not any(p.requires_grad for p in self.parameters()):
            return self.forward(x, pose)
        return _FusedMLPFunction.apply(self, x, pose, *self.parameters())


class _FusedMLPFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mlp: FusedSkipMLP, x, pose, *params):
        out, state = mlp.forward(x, pose, save=True)
        ctx.mlp, ctx.state = mlp, state
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, g_out):
        (out,) = ctx.saved_tensors
        grads = ctx.mlp.backward(g_out, out, ctx.state)
        ctx.state = None
        return (None, None, None, *grads)
