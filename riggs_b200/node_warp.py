"""Stage-1 control-node warp on the kernels of csrc/node_warp.cu: host-side mirror of
``utils.time_utils.ControlNodeWarp`` (utils/time_utils.py:770-1236) for the per-frame path -- ``cal_nn_weight``
(:934-967, K nearest nodes + radial-basis weights) and the blending of ``forward`` (:1133-1236) -- with the same
names, argument meaning and return dicts.  The node-motion network (``node_deform``: an 8x256 MLP, dense GEMMs)
stays a torch module supplied by the caller; node initialisation (farthest-point sampling), densification, the
ARAP regulariser and the GUI's drag-editing branch (``node_trans_bias``) are not part of the
hot path and are not mirrored.  No CPU path.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch
from torch import nn

from . import _lib
from ._lib import check, f32c, ptr, stream_ptr

# columns of the per-node gradient table written by rg_node_warp_backward (include/riggs_b200.h)
_COLS = {"trans": (0, 3), "rot": (3, 7), "scale": (7, 10), "quat": (10, 14), "radius": (14, 15), "weight": (15, 16),
         "opacity": (16, 17), "color": (17, 20), "hyper": (20, 25)}


class _NodeWarp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, feature, nodes, radius_raw, weight_raw, local_quat, node_trans, node_rot, node_scale,
                node_opacity, node_color, mask, K, hyper, add_rot_bias):
        if not x.is_cuda:
            raise _lib.RiggsError("the control-node warp needs CUDA tensors; riggs_b200 has no CPU path")
        N, M = x.shape[0], nodes.shape[0]
        dev = x.device
        ctx.set_materialize_grads(False)
        x_, nodes_ = f32c(x[:, :3].detach()), f32c(nodes.detach()[:, :3 + hyper])
        feat_ = f32c(feature.detach()[:, :hyper]) if hyper > 0 else None
        t = dict(radius=f32c(radius_raw.detach().reshape(-1)),
                 weight=f32c(weight_raw.detach().reshape(-1)) if weight_raw is not None else None,
                 quat=f32c(local_quat.detach()) if local_quat is not None else None,
                 trans=f32c(node_trans.detach()), rot=f32c(node_rot.detach()), scale=f32c(node_scale.detach()),
                 opacity=f32c(node_opacity.detach().reshape(-1)) if node_opacity is not None else None,
                 color=f32c(node_color.detach()) if node_color is not None else None,
                 mask=f32c(mask.detach().reshape(-1)) if mask is not None else None)
        d_xyz = torch.empty(N, 3, device=dev)
        d_rot = torch.empty(N, 4, device=dev)
        d_scale = torch.empty(N, 3, device=dev)
        d_op = torch.empty(N, 1, device=dev) if node_opacity is not None else None
        d_col = torch.empty(N, 3, device=dev) if node_color is not None else None
        nn_idx = torch.empty(N, K, dtype=torch.int64, device=dev)
        nn_w = torch.empty(N, K, device=dev)
        nn_d = torch.empty(N, K, device=dev)
        with torch.cuda.device(dev):
            check(_lib.lib().rg_node_warp_forward(
                N, M, K, hyper, ptr(x_), ptr(feat_), ptr(nodes_), ptr(t["radius"]), ptr(t["weight"]), ptr(t["quat"]),
                ptr(t["trans"]), ptr(t["rot"]), ptr(t["scale"]), ptr(t["opacity"]), ptr(t["color"]), ptr(t["mask"]),
                int(add_rot_bias), ptr(d_xyz), ptr(d_rot), ptr(d_scale), ptr(d_op), ptr(d_col), ptr(nn_idx), ptr(nn_w),
                ptr(nn_d), stream_ptr()), "rg_node_warp_forward")
        ctx.save_for_backward(x_, feat_, nodes_, nn_idx, nn_d, *[v for v in t.values() if v is not None])
        ctx.present = [k for k, v in t.items() if v is not None]
        ctx.dims = (N, M, K, hyper, nodes.shape[1], feature.shape if feature is not None else None,
                    tuple(radius_raw.shape), tuple(weight_raw.shape) if weight_raw is not None else None,
                    tuple(node_opacity.shape) if node_opacity is not None else None,
                    tuple(mask.shape) if mask is not None else None)
        ctx.mark_non_differentiable(nn_idx)
        return d_xyz, d_rot, d_scale, d_op, d_col, nn_idx, nn_w, nn_d

    @staticmethod
    def backward(ctx, g_xyz, g_rot, g_scale, g_op, g_col, _g_idx, g_w, g_d):
        if g_w is not None or g_d is not None:
            # nn_weight / nn_dist are differentiable in the reference; its trainers never take their gradient
            raise _lib.RiggsError("the control-node warp has no backward for nn_weight / nn_dist cotangents")
        x_, feat_, nodes_, nn_idx, nn_d, *rest = ctx.saved_tensors
        t = dict(zip(ctx.present, rest))
        N, M, K, hyper, node_cols, feat_shape, radius_shape, weight_shape, op_shape, mask_shape = ctx.dims
        dev = x_.device
        L = _lib.lib()
        C = int(L.rg_node_warp_grad_channels())
        dnode = torch.zeros(M, C, device=dev)
        need = ctx.needs_input_grad
        dmask = torch.empty(N, device=dev) if (mask_shape is not None and need[11]) else None
        dfeat = torch.empty(N, hyper, device=dev) if (hyper > 0 and need[1]) else None
        g = lambda v: f32c(v) if v is not None else None
        g_xyz, g_rot, g_scale, g_op, g_col = g(g_xyz), g(g_rot), g(g_scale), g(g_op), g(g_col)
        with torch.cuda.device(dev):
            check(L.rg_node_warp_backward(
                N, M, K, hyper, ptr(x_), ptr(feat_), ptr(nodes_), ptr(t["radius"]), ptr(t.get("weight")),
                ptr(t.get("quat")), ptr(t["trans"]), ptr(t["rot"]), ptr(t["scale"]), ptr(t.get("opacity")),
                ptr(t.get("color")), ptr(t.get("mask")), ptr(nn_idx), ptr(nn_d), ptr(g_xyz), ptr(g_rot), ptr(g_scale),
                ptr(g_op), ptr(g_col), ptr(dnode), ptr(dmask), ptr(dfeat), stream_ptr()), "rg_node_warp_backward")
        col = lambda k: dnode[:, _COLS[k][0]:_COLS[k][1]]
        d_nodes = None
        if need[2] and hyper > 0:
            d_nodes = torch.zeros(M, node_cols, device=dev)
            d_nodes[:, 3:3 + hyper] = dnode[:, 20:20 + hyper]
        d_feature = None
        if dfeat is not None:
            d_feature = torch.zeros(feat_shape, device=dev)
            d_feature[:, :hyper] = dfeat
        return (None, d_feature, d_nodes, col("radius").reshape(radius_shape),
                col("weight").reshape(weight_shape) if weight_shape is not None else None,
                col("quat") if "quat" in t else None, col("trans"), col("rot"), col("scale"),
                col("opacity").reshape(op_shape) if op_shape is not None else None,
                col("color") if "color" in t else None,
                dmask.reshape(mask_shape) if dmask is not None else None, None, None, None)


def node_warp(x, feature, nodes, node_radius_raw, node_weight_raw, node_attrs: Dict[str, torch.Tensor], motion_mask,
              K: int = 3, hyper_dim: int = 2, local_frame: bool = True, d_rot_as_res: bool = True) -> Dict[str, object]:
    """Functional form: one forward launch (one backward launch under autograd).  ``node_attrs`` as produced by the
    reference's ``node_deform``: ``d_xyz (M,3), d_rotation (M,4), d_scaling (M,3)`` and optionally
    ``local_rotation (M,4), d_opacity (M,1), d_color (M,3)``."""
    hyper = hyper_dim if (hyper_dim > 0 and feature is not None) else 0
    quat = None
    if local_frame:
        quat = node_attrs["local_rotation"] + torch.tensor([1.0, 0.0, 0.0, 0.0], device=x.device)
    mask = motion_mask if torch.is_tensor(motion_mask) else None
    scale_by = None if mask is not None or motion_mask is None else float(motion_mask)
    out = _NodeWarp.apply(x, feature, nodes, node_radius_raw, node_weight_raw, quat, node_attrs["d_xyz"],
                          node_attrs["d_rotation"], node_attrs["d_scaling"], node_attrs.get("d_opacity"),
                          node_attrs.get("d_color"), mask, int(K), int(hyper), not d_rot_as_res)
    d_xyz, d_rot, d_scale, d_op, d_col, nn_idx, nn_w, nn_d = out
    if scale_by is not None and scale_by != 1.0:  # a python scalar mask (the GUI passes floats)
        bias = torch.tensor([1.0, 0, 0, 0], device=x.device) if not d_rot_as_res else 0.0
        d_xyz, d_scale = d_xyz * scale_by, d_scale * scale_by
        d_rot = (d_rot - bias) * scale_by + bias
        d_op = d_op * scale_by if d_op is not None else None
        d_col = d_col * scale_by if d_col is not None else None
    return {"d_xyz": d_xyz, "d_rotation": d_rot, "d_scaling": d_scale, "d_nodes": nodes[..., :3] + node_attrs["d_xyz"],
            "d_opacity": d_op, "d_color": d_col, "nn_idx": nn_idx, "nn_weight": nn_w, "nn_dist": nn_d}


class ControlNodeWarp(nn.Module):
    """Parameters and per-frame methods of the reference module (utils/time_utils.py:770-967, 1133-1236).

    ``network``: callable ``t (M,1) -> dict`` with the keys of the reference's ``node_deform`` (a torch module whose
    parameters are trained with the nodes); ``forward`` without a network needs ``node_attrs``."""

    def __init__(self, nodes: torch.Tensor, node_radius_raw: torch.Tensor, node_weight_raw: Optional[torch.Tensor] = None,
                 K: int = 3, hyper_dim: int = 2, local_frame: bool = False, d_rot_as_res: bool = True,
                 with_node_weight: bool = True, pred_opacity: bool = False, pred_color: bool = False,
                 network: Optional[Callable] = None, device="cuda"):
        super().__init__()
        self.K, self.hyper_dim, self.local_frame, self.d_rot_as_res = K, hyper_dim, local_frame, d_rot_as_res
        self.with_node_weight, self.pred_opacity, self.pred_color = with_node_weight, pred_opacity, pred_color
        self.skinning = False
        self.name = "node"
        self.nodes = nn.Parameter(nodes.float().to(device))
        self._node_radius = nn.Parameter(node_radius_raw.float().to(device))
        if with_node_weight:
            w = node_weight_raw if node_weight_raw is not None else torch.zeros(nodes.shape[0], 1)
            self._node_weight = nn.Parameter(w.float().to(device))
        self.network = network
        self.cached_nn_weight = False
        self.nn_weight = self.nn_dist = self.nn_idxs = None
        self.reg_loss = 0.0

    @property
    def node_radius(self):
        return torch.exp(self._node_radius)

    @property
    def node_weight(self):
        return torch.sigmoid(self._node_weight)

    @property
    def node_num(self):
        return self.nodes.shape[0]

    def expand_time(self, t):
        return t.unsqueeze(0).expand(self.nodes.shape[0], -1)

    def node_deform(self, t, **kwargs):
        if self.network is None:
            raise _lib.RiggsError("ControlNodeWarp has no node-motion network; pass node_attrs")
        return self.network(t, **kwargs)

    def _zero_attrs(self):
        M, dev = self.node_num, self.nodes.device
        return {"d_xyz": torch.zeros(M, 3, device=dev), "d_rotation": torch.zeros(M, 4, device=dev),
                "d_scaling": torch.zeros(M, 3, device=dev), "local_rotation": torch.zeros(M, 4, device=dev)}

    def cal_nn_weight(self, x: torch.Tensor, K=None, feature=None, nodes=None, gs_kernel=True, temperature=1.0):
        """(nn_weight (N,K), nn_dist (N,K), nn_idx (N,K) int64) as the reference returns them (:934-967)."""
        if not gs_kernel:
            raise NotImplementedError("cal_nn_weight(gs_kernel=False) is only used by the animation branches")
        if self.cached_nn_weight and self.nn_weight is not None:
            return self.nn_weight, self.nn_dist, self.nn_idxs
        nd = self.nodes if nodes is None else nodes
        attrs = self._zero_attrs()
        if nd.shape[0] != self.node_num:
            raise ValueError("cal_nn_weight: nodes must have one row per control node")
        out = node_warp(x, feature, nd, self._node_radius, self._node_weight if self.with_node_weight else None, attrs,
                        None, K=self.K if K is None else K, hyper_dim=self.hyper_dim, local_frame=False)
        if self.cached_nn_weight:
            self.nn_weight, self.nn_dist, self.nn_idxs = out["nn_weight"], out["nn_dist"], out["nn_idx"]
        return out["nn_weight"], out["nn_dist"], out["nn_idx"]

    def forward(self, x, t, feature, motion_mask, iteration=0, is_training=True, node_attrs=None, node_trans_bias=None,
                node_scaling_bias=None, animation_d_values=None, **kwargs):
        """``animation_d_values``: per-key overrides of the node attributes (utils/time_utils.py:1142-1144: a motion
        editor replaces the network's ``d_xyz`` / ``d_rotation`` / ... by its own values).  ``node_trans_bias`` -- the GUI's
        drag editing (:1165-1204: per-node rotations by SVD over trajectories, geodesic re-weighting) -- belongs to
        ``train_gui.py`` and is not part of this path."""
        if node_trans_bias is not None:
            raise NotImplementedError("ControlNodeWarp.forward(node_trans_bias=...) is the GUI's drag editing "
                                      "(train_gui.py); not part of the hot path")
        if node_attrs is None:
            if t.dim() == 0:
                t = self.expand_time(t)
            node_attrs = self.node_deform(t=t, **kwargs)
        attrs = dict(node_attrs)
        if animation_d_values is not None:
            for key in animation_d_values:
                attrs[key] = animation_d_values[key]
        if not self.pred_opacity:
            attrs.pop("d_opacity", None)
        if not self.pred_color:
            attrs.pop("d_color", None)
        out = node_warp(x.detach(), feature, self.nodes, self._node_radius,
                        self._node_weight if self.with_node_weight else None, attrs, motion_mask, K=self.K,
                        hyper_dim=self.hyper_dim, local_frame=self.local_frame, d_rot_as_res=self.d_rot_as_res)
        self.reg_loss = 0.0
        return {k: out[k] for k in ("d_xyz", "d_rotation", "d_scaling", "d_nodes", "d_opacity", "d_color")}

    def trainable_parameters(self):
        group = [self.nodes, self._node_radius] + ([self._node_weight] if self.with_node_weight else [])
        params = [{"params": group, "name": "nodes"}]
        if isinstance(self.network, nn.Module):
            params.insert(0, {"params": list(self.network.parameters()), "name": "deform"})
        return params
