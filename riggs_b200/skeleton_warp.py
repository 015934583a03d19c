"""Host mirror of the reference's skeleton skinning interface for the hot path.

``SkeletonWarp.forward(x, t, motion_mask)`` and ``SkeletonWarp.deform_by_pose(x, node_attrs, motion_mask)``
keep the names, argument meaning and the returned dict of skeleton_utils/skeleton_warp.py:112-172
(keys ``d_xyz, d_rotation, d_scaling, d_nodes, nn_idx, nn_weight, local_rotation, global_trans, d_opacity,
d_color``); ``SkeletonModel.step`` keeps scene/skeleton_model.py:21-22.  The LBS core and the forward
kinematics run as CUDA kernels of libriggs_b200.so with hand-written backward kernels; the three small MLP
front ends (PoseMLP / WeightMLP / DeformMLP, skeleton_utils/network_utils.py) stay plain torch modules --
they are dense GEMMs and out of this path's scope.

Gradient contract (skeleton_warp.py:131,144): nothing flows to ``x`` or through ``node_rot``; gradients reach
``local_rotation`` (through the FK chain), ``global_trans``, ``_node_radius``, ``motion_mask``, the WeightMLP
output and the detail-net offsets.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from ._lib import check, f32c, ptr, stream_ptr


class _FKFunction(torch.autograd.Function):
    """(F,J,4) local quaternions, (F,3) translations -> G (F,J,12), node_rot (F,J,4), d_nodes (F,J,3)."""

    @staticmethod
    def forward(ctx, local_rot, global_trans, joints, parents_i32):
        if not local_rot.is_cuda:
            raise _lib.RiggsError("forward kinematics needs CUDA tensors; riggs_b200 has no CPU path")
        local_rot, global_trans, joints = f32c(local_rot), f32c(global_trans), f32c(joints)
        Fn, J = local_rot.shape[0], local_rot.shape[1]
        dev = local_rot.device
        G = torch.empty((Fn, J, 12), dtype=torch.float32, device=dev)
        node_rot = torch.empty((Fn, J, 4), dtype=torch.float32, device=dev)
        posed = torch.empty((Fn, J, 3), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().rg_fk_forward(Fn, J, ptr(local_rot), ptr(global_trans), ptr(joints), ptr(parents_i32),
                                           ptr(G), ptr(node_rot), ptr(posed), stream_ptr()), "fk_forward")
        ctx.save_for_backward(local_rot, joints, parents_i32, G)
        ctx.mark_non_differentiable(node_rot)
        return G, node_rot, posed

    @staticmethod
    def backward(ctx, dG, _dnode_rot, dposed):
        local_rot, joints, parents_i32, G = ctx.saved_tensors
        Fn, J = local_rot.shape[0], local_rot.shape[1]
        dev = local_rot.device
        dG = f32c(dG) if dG is not None else None
        dposed = f32c(dposed) if dposed is not None else None
        dq = torch.empty((Fn, J, 4), dtype=torch.float32, device=dev)
        dtr = torch.zeros((Fn, 3), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().rg_fk_backward(Fn, J, ptr(local_rot), ptr(joints), ptr(parents_i32), ptr(G), ptr(dG),
                                            ptr(dposed), ptr(dq), ptr(dtr), stream_ptr()), "fk_backward")
        return dq, dtr, None, None


class _SkinFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask, mlp_weight,
                offsets, blend_mode, want_weights):
        if not xyz.is_cuda:
            raise _lib.RiggsError("skinning needs CUDA tensors; riggs_b200 has no CPU path")
        xyz, joints, node_radius_raw, G, node_rot, global_trans = (f32c(t) for t in (
            xyz, joints, node_radius_raw, G, node_rot, global_trans))
        motion_mask, mlp_weight, offsets = f32c(motion_mask), f32c(mlp_weight), f32c(offsets)
        N, J = xyz.shape[0], joints.shape[0]
        dev = xyz.device
        d_xyz = torch.empty((N, 3), dtype=torch.float32, device=dev)
        d_rot = torch.empty((N, 4), dtype=torch.float32, device=dev)
        nn_w = torch.empty((N, J - 1), dtype=torch.float32, device=dev) if want_weights else None
        with torch.cuda.device(dev):
            check(_lib.lib().rg_skin_forward(N, J, ptr(xyz), ptr(joints), ptr(parents_i32), ptr(node_radius_raw),
                                             ptr(G), ptr(node_rot), ptr(global_trans), ptr(motion_mask),
                                             ptr(mlp_weight), ptr(offsets), int(blend_mode), ptr(d_xyz), ptr(d_rot),
                                             ptr(nn_w), stream_ptr()), "skin_forward")
        ctx.save_for_backward(xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask,
                              mlp_weight, offsets)
        ctx.blend_mode = blend_mode
        if nn_w is None:
            nn_w = torch.empty(0, device=dev)
        ctx.mark_non_differentiable(nn_w)
        return d_xyz, d_rot, nn_w

    @staticmethod
    def backward(ctx, g_dxyz, g_drot, _g_w):
        (xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask, mlp_weight,
         offsets) = ctx.saved_tensors
        N, J = xyz.shape[0], joints.shape[0]
        dev = xyz.device
        g_dxyz = f32c(g_dxyz) if g_dxyz is not None else torch.zeros((N, 3), dtype=torch.float32, device=dev)
        g_drot = f32c(g_drot) if g_drot is not None else None
        dG = torch.zeros((J, 12), dtype=torch.float32, device=dev)
        dradius = torch.zeros((J,), dtype=torch.float32, device=dev)
        dtrans = torch.zeros((3,), dtype=torch.float32, device=dev)
        need = ctx.needs_input_grad
        dmask = torch.empty_like(motion_mask) if (motion_mask is not None and need[7]) else None
        dmlp = torch.empty_like(mlp_weight) if (mlp_weight is not None and need[8]) else None
        doff = torch.empty_like(offsets) if (offsets is not None and need[9]) else None
        with torch.cuda.device(dev):
            check(_lib.lib().rg_skin_backward_blend(N, J, ptr(xyz), ptr(joints), ptr(parents_i32), ptr(node_radius_raw),
                                                    ptr(G), ptr(node_rot), ptr(global_trans), ptr(motion_mask),
                                                    ptr(mlp_weight), ptr(offsets), int(ctx.blend_mode), ptr(g_dxyz),
                                                    ptr(g_drot), ptr(dG), ptr(dradius), ptr(dtrans), ptr(dmask), ptr(dmlp),
                                                    ptr(doff), stream_ptr()), "skin_backward")
        return (None, None, None, dradius, dG.view_as(G), None, dtrans.view_as(global_trans), dmask, dmlp, doff,
                None, None)


class _SkinTopKFunction(torch.autograd.Function):
    """LBS by the K nearest bones (cal_nn_weight_skeleton with K > 0, skeleton_warp.py:46-49) on csrc/skin_topk.cu."""

    @staticmethod
    def forward(ctx, xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask, offsets, K):
        if not xyz.is_cuda:
            raise _lib.RiggsError("skinning needs CUDA tensors; riggs_b200 has no CPU path")
        ctx.set_materialize_grads(False)
        xyz, joints, node_radius_raw, G, node_rot, global_trans = (f32c(t) for t in (
            xyz, joints, node_radius_raw, G, node_rot, global_trans))
        motion_mask, offsets = f32c(motion_mask), f32c(offsets)
        N, J = xyz.shape[0], joints.shape[0]
        dev = xyz.device
        d_xyz = torch.empty((N, 3), dtype=torch.float32, device=dev)
        d_rot = torch.empty((N, 4), dtype=torch.float32, device=dev)
        nn_w = torch.empty((N, K), dtype=torch.float32, device=dev)
        nn_idx = torch.empty((N, K), dtype=torch.int64, device=dev)
        nn_d2 = torch.empty((N, K), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().rg_skin_topk_forward(N, J, int(K), ptr(xyz), ptr(joints), ptr(parents_i32),
                                                  ptr(node_radius_raw), ptr(G), ptr(node_rot), ptr(global_trans),
                                                  ptr(motion_mask), ptr(offsets), ptr(d_xyz), ptr(d_rot), ptr(nn_w),
                                                  ptr(nn_idx), ptr(nn_d2), stream_ptr()), "rg_skin_topk_forward")
        ctx.save_for_backward(xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask, offsets,
                              nn_idx)
        ctx.K = int(K)
        ctx.mark_non_differentiable(nn_w, nn_idx, nn_d2)
        return d_xyz, d_rot, nn_w, nn_idx, nn_d2

    @staticmethod
    def backward(ctx, g_dxyz, g_drot, _gw, _gi, _gd):
        (xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask, offsets,
         nn_idx) = ctx.saved_tensors
        N, J = xyz.shape[0], joints.shape[0]
        dev = xyz.device
        g_dxyz = f32c(g_dxyz) if g_dxyz is not None else None
        g_drot = f32c(g_drot) if g_drot is not None else None
        dG = torch.zeros((J, 12), dtype=torch.float32, device=dev)
        dradius = torch.zeros((J,), dtype=torch.float32, device=dev)
        dtrans = torch.zeros((3,), dtype=torch.float32, device=dev)
        need = ctx.needs_input_grad
        dmask = torch.empty_like(motion_mask) if (motion_mask is not None and need[7]) else None
        doff = torch.empty_like(offsets) if (offsets is not None and need[8]) else None
        with torch.cuda.device(dev):
            check(_lib.lib().rg_skin_topk_backward(N, J, ctx.K, ptr(xyz), ptr(joints), ptr(parents_i32),
                                                   ptr(node_radius_raw), ptr(G), ptr(node_rot), ptr(global_trans),
                                                   ptr(motion_mask), ptr(offsets), ptr(nn_idx), ptr(g_dxyz), ptr(g_drot),
                                                   ptr(dG), ptr(dradius), ptr(dtrans), ptr(dmask), ptr(doff), stream_ptr()),
                  "rg_skin_topk_backward")
        return None, None, None, dradius, dG.view_as(G), None, dtrans.view_as(global_trans), dmask, doff, None


def skin_topk(xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask=None, offsets=None, K=3):
    """(d_xyz, d_rotation, nn_weight (N,K), nn_idx (N,K) int64 joint numbers, nn_dist2 (N,K))."""
    return _SkinTopKFunction.apply(xyz.detach(), joints, parents_i32, node_radius_raw, G, node_rot, global_trans,
                                   motion_mask, offsets, int(K))


def fk(local_rot: torch.Tensor, global_trans: torch.Tensor, joints: torch.Tensor, parents_i32: torch.Tensor):
    """Batched forward kinematics.  local_rot (F,J,4) or (J,4); returns G, node_rot, d_nodes with a leading F."""
    single = local_rot.dim() == 2
    if single:
        local_rot, global_trans = local_rot[None], global_trans[None]
    G, node_rot, posed = _FKFunction.apply(local_rot, global_trans, joints, parents_i32)
    if single:
        return G[0], node_rot[0], posed[0]
    return G, node_rot, posed


def skin(xyz, joints, parents_i32, node_radius_raw, G, node_rot, global_trans, motion_mask=None, mlp_weight=None,
         offsets=None, blend_mode=0, want_weights=False):
    d_xyz, d_rot, w = _SkinFunction.apply(xyz.detach(), joints, parents_i32, node_radius_raw, G, node_rot,
                                          global_trans, motion_mask, mlp_weight, offsets, blend_mode, want_weights)
    return d_xyz, d_rot, (w if want_weights else None)


# ---- MLP front ends kept in torch (skeleton_utils/network_utils.py) -------------------------------------

class _Embedder(nn.Module):
    """NeRF positional encoding with include_input (utils/time_utils.py:208-256)."""

    def __init__(self, multires, dims):
        super().__init__()
        self.freqs = [2.0 ** i for i in range(multires)]
        self.out_dim = dims * (1 + 2 * multires)

    def forward(self, x):
        out = [x]
        for f in self.freqs:
            out += [torch.sin(x * f), torch.cos(x * f)]
        return torch.cat(out, -1)


def _skip_layers(in_ch, width=256, depth=8):
    """The 8 x 256 ReLU stack with the embedded input re-fed after layer depth // 2 (network_utils.py:28-32, 85-89,
    127-128).  A plain ModuleList, so that its parameters are named ``<attr>.{i}.weight`` like the reference's."""
    skips = [depth // 2]
    return nn.ModuleList([nn.Linear(in_ch, width)] + [
        nn.Linear(width, width) if i not in skips else nn.Linear(width + in_ch, width) for i in range(depth - 1)])


def _skip_forward(layers, e, skip=4):
    h = e
    for i, l in enumerate(layers):
        h = F.relu(l(h))
        if i == skip:
            h = torch.cat([e, h], -1)
    return h


class PoseMLP(nn.Module):
    """t -> {'rotation': (4J,), 'translation': (3,)} (network_utils.py:115-150).  Parameter names are the reference's
    (``net.{i}``, ``rotation_predictor``, ``translation_predictor``), so a ``skeleton.pth`` loads by key."""

    def __init__(self, input_ch, output_ch, multires=8):
        super().__init__()
        self.embed = _Embedder(multires, input_ch)
        self.net = _skip_layers(self.embed.out_dim)
        self.rotation_predictor = nn.Linear(256, output_ch)
        self.translation_predictor = nn.Linear(256, 3)

    def forward(self, t):
        h = _skip_forward(self.net, self.embed(t))
        return {"rotation": self.rotation_predictor(h), "translation": self.translation_predictor(h)}


class WeightMLP(nn.Module):
    """x -> sigmoid (J-1) per-bone weight modulation (network_utils.py:73-107; parameters ``linear.{i}``,
    ``weight_predict``)."""

    def __init__(self, input_ch, output_ch, multires=10):
        super().__init__()
        self.embed = _Embedder(multires, 3)
        self.linear = _skip_layers(self.embed.out_dim)
        self.weight_predict = nn.Linear(256, output_ch)

    def forward(self, x):
        return torch.sigmoid(self.weight_predict(_skip_forward(self.linear, self.embed(x))))


class DeformMLP(nn.Module):
    """(x, pose) -> 3-D template offset (network_utils.py:6-65; parameters ``linear.{i}``, ``gaussian_warp``; the stack is
    He-initialised with zero biases and the head starts at N(0, 1e-5) like the reference's, :35-40)."""

    def __init__(self, time_input_ch, multires=4):
        super().__init__()
        self.embed = _Embedder(multires, 3)
        self.linear = _skip_layers(self.embed.out_dim + time_input_ch)
        self.gaussian_warp = nn.Linear(256, 3)
        for layer in self.linear:
            nn.init.kaiming_uniform_(layer.weight, mode="fan_in", nonlinearity="relu")
            nn.init.zeros_(layer.bias)
        nn.init.normal_(self.gaussian_warp.weight, mean=0, std=1e-5)
        nn.init.zeros_(self.gaussian_warp.bias)

    def forward(self, x, t):
        return self.gaussian_warp(_skip_forward(self.linear, torch.cat([self.embed(x), t], -1)))


class SkeletonWarp(nn.Module):
    """CUDA-backed counterpart of skeleton_utils/skeleton_warp.py:10-172 (K = -1: all bones; K > 0: the K nearest)."""

    def __init__(self, joints: torch.Tensor, parent_indices: torch.Tensor, node_radius_raw: Optional[torch.Tensor] = None,
                 hyper_dim: int = 2, use_skinning_weight_mlp: bool = True, use_template_offsets: bool = True,
                 blend_mode: str = "lbs", device="cuda", K: int = -1):
        super().__init__()
        J = joints.shape[0]
        nodes = torch.cat([joints.float(), 1e-2 * torch.ones(J, hyper_dim)], -1)
        self.nodes = nn.Parameter(nodes.to(device), requires_grad=False)
        self.register_buffer("parents", parent_indices.long().to(device))
        self.register_buffer("parents_i32", parent_indices.to(torch.int32).to(device))
        if node_radius_raw is None:
            rng = joints.max() - joints.min()
            node_radius_raw = torch.log(0.1 * rng + 1e-7) * torch.ones(J)
        self._node_radius = nn.Parameter(node_radius_raw.float().to(device))
        self.use_skinning_weight_mlp = use_skinning_weight_mlp
        self.use_template_offsets = use_template_offsets
        self.blend_mode = {"lbs": 0, "dq": 1}[blend_mode]
        self.skinning_weight_mlp = WeightMLP(3, J - 1).to(device) if use_skinning_weight_mlp else None
        self.detail_net = DeformMLP(J * 4).to(device) if use_template_offsets else None
        self.pose_net = PoseMLP(1, J * 4).to(device)
        self.skinning_weight_offsets = None
        self.template_offsets = None
        self.cache_skinning_weights = False  # see _weight_mlp_output
        self._weight_cache = None
        self.K = int(K)  # -1: all bones (what the reference's scripts run); > 0: the K nearest bones

    @property
    def node_radius(self):
        return torch.exp(self._node_radius)

    def forward(self, x, t, motion_mask, **kwargs):
        if t.dim() == 0:
            t = t.unsqueeze(0).expand(self.nodes.shape[0], -1)
        rot_bias = torch.tensor([1., 0, 0, 0], device=x.device)
        motion = self.pose_net(t[0])
        node_attrs = {"local_rotation": motion["rotation"].reshape(-1, 4) + rot_bias,
                      "global_trans": motion["translation"], "t": t[0]}
        return self.deform_by_pose(x, node_attrs, motion_mask)

    def deform_by_pose(self, x, node_attrs, motion_mask, want_weights: bool = True):
        x = x.detach()
        local_rot, global_trans = node_attrs["local_rotation"], node_attrs["global_trans"]
        joints = self.nodes[:, :3].detach()
        G, node_rot, d_nodes = fk(local_rot, global_trans, joints, self.parents_i32)
        mlp_w = None
        if self.use_skinning_weight_mlp and self.skinning_weight_mlp is not None:
            mlp_w = self._weight_mlp_output(x)
            self.skinning_weight_offsets = mlp_w
        offs = None
        if self.use_template_offsets and self.detail_net is not None:
            pose = local_rot.detach().reshape(-1)[None].expand(x.shape[0], -1)
            offs = self.detail_net(x, pose)
        self.template_offsets = offs if offs is not None else torch.zeros_like(x)
        mask = motion_mask.reshape(-1) if torch.is_tensor(motion_mask) else None
        J = joints.shape[0]
        if self.K > 0:
            if mlp_w is not None:
                # the reference gathers the MLP output by joint number (skeleton_warp.py:59), which is out of range for
                # the last bone: there is no defined behaviour to mirror
                raise _lib.RiggsError("SkeletonWarp with K > 0 needs use_skinning_weight_mlp = False")
            if self.blend_mode != 0:
                raise _lib.RiggsError("SkeletonWarp with K > 0 supports linear blending only")
            d_xyz, d_rot, w, nn_idx, _ = skin_topk(x, joints, self.parents_i32, self._node_radius, G, node_rot,
                                                   global_trans, mask, offs, self.K)
        else:
            d_xyz, d_rot, w = skin(x, joints, self.parents_i32, self._node_radius, G, node_rot, global_trans, mask, mlp_w,
                                   offs, self.blend_mode, want_weights)
            nn_idx = torch.arange(1, J, dtype=torch.long, device=x.device).unsqueeze(0).expand(x.shape[0], -1)
        return {"d_xyz": d_xyz, "d_rotation": d_rot, "d_scaling": torch.zeros(x.shape[0], 3, device=x.device),
                "d_nodes": d_nodes, "nn_idx": nn_idx, "nn_weight": w, "local_rotation": local_rot,
                "global_trans": global_trans, "d_opacity": None, "d_color": None}

    def _weight_mlp_output(self, x):
        """sigmoid(WeightMLP(x)) (skeleton_warp.py:56-61).  It depends on the canonical positions only, not on the pose,
        so with ``cache_skinning_weights = True`` (novel-pose rendering of a frozen model: the 8 x 256 MLP over every
        Gaussian costs far more than skinning and rasterizing a frame) it is evaluated once per canonical cloud and
        reused for every frame; the cache is keyed on the tensor's storage, shape and in-place version and is bypassed
        whenever autograd is recording."""
        if not self.cache_skinning_weights or torch.is_grad_enabled():
            return self.skinning_weight_mlp(x[..., :3])
        key = (x.data_ptr(), tuple(x.shape), x._version, str(x.device))
        if self._weight_cache is None or self._weight_cache[0] != key:
            self._weight_cache = (key, self.skinning_weight_mlp(x[..., :3]).detach())
        return self._weight_cache[1]

    def clear_weight_cache(self):
        """Call after changing the weight MLP's parameters (e.g. loading a checkpoint) while caching is on."""
        self._weight_cache = None

    def trainable_parameters(self):
        params = [{"params": [self._node_radius], "name": "nodes"},
                  {"params": list(self.pose_net.parameters()), "name": "pose"}]
        if self.use_skinning_weight_mlp and self.skinning_weight_mlp is not None:
            params.append({"params": list(self.skinning_weight_mlp.parameters()), "name": "skinning_mlp"})
        if self.use_template_offsets and self.detail_net is not None:
            params.append({"params": list(self.detail_net.parameters()), "name": "detail_net"})
        return params


class SkeletonModel:
    """scene/skeleton_model.py:8-22: owner of the warp module; ``step`` delegates to it."""

    def __init__(self, joints, parent_indices, **kw):
        self.deform = SkeletonWarp(joints, parent_indices, **kw)

    def step(self, x, time_emb, **kwargs):
        return self.deform(x, time_emb, **kwargs)
