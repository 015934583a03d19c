"""One iteration of the hot path as the reference's trainer runs it (train_rig.py:535-554 minus SSIM and the
regularisers): pose -> forward kinematics -> skin all Gaussians -> render one view -> L1 loss -> backward to
every Gaussian parameter and to the skeleton inputs.  Used by bench.py, the smoke test and the multi-GPU path.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import skeleton_warp as sw
from .gaussian_renderer import GaussianParams, render


class _Pipe:
    debug = False
    convert_SHs_python = False
    compute_cov3D_python = False


class HotPath:
    def __init__(self, scene, device="cuda", use_mlp_weight: bool = False, use_offsets: bool = False,
                 fused: bool = True):
        """fused=True: one C-ABI call per direction (rg_skinned_forward/backward, no host sync);
        fused=False: the reference's call sequence (deform_by_pose, then render()) on the same kernels."""
        self.fused = fused
        self.device = torch.device(device)
        self.pc = GaussianParams.from_scene(scene, device=device)
        self.warp = sw.SkeletonWarp(scene.joints, scene.parents, scene.node_radius_raw,
                                    use_skinning_weight_mlp=False, use_template_offsets=False, device=device)
        self.mlp_weight = scene.mlp_weight.to(device) if (use_mlp_weight and scene.mlp_weight is not None) else None
        self.offsets = scene.template_offsets.to(device) if (use_offsets and scene.template_offsets is not None) else None
        self.pipe = _Pipe()
        self.bg = torch.zeros(3, device=device)
        self.isotropic = self.pc.use_isotropic_gs
        self.flat = None

    def named_parameters(self):
        pc = self.pc
        return [("xyz", pc._xyz), ("features_dc", pc._features_dc), ("features_rest", pc._features_rest),
                ("opacity", pc._opacity), ("scaling", pc._scaling), ("rotation", pc._rotation),
                ("node_radius", self.warp._node_radius)]

    def enable_flat_grads(self):
        """Data-parallel mode: the fused backward writes the parameter gradients straight into one flat buffer
        (dist.FlatGrads) that is all-reduced in place; ``reduce_grads()`` then exposes it through ``param.grad``."""
        from .dist import FlatGrads
        self.flat = FlatGrads(self.named_parameters(), device=self.device)
        return self.flat

    def reduce_grads(self):
        if self.flat is None:
            return
        if not self.fused:
            self.flat.pack_from_grads()
        self.flat.all_reduce()
        self.flat.attach()

    def parameters(self):
        pc = self.pc
        return [pc._xyz, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation,
                self.warp._node_radius]

    def zero_grad(self):
        for p in self.parameters():
            p.grad = None

    def forward(self, cam, pose: Dict[str, torch.Tensor]):
        """pose: {'local_rotation': (J,4), 'global_trans': (3,)} on the device (may require grad)."""
        if self.fused:
            from .fused import render_skinned
            return render_skinned(cam, self.pc, self.warp, pose, self.pipe, self.bg, mlp_weight=self.mlp_weight,
                                  offsets=self.offsets, grad_out=self.flat.views if self.flat is not None else None)
        joints = self.warp.nodes[:, :3]
        G, node_rot, d_nodes = sw.fk(pose["local_rotation"], pose["global_trans"], joints, self.warp.parents_i32)
        mask = self.pc._motion_mask.reshape(-1) if self.pc._motion_mask is not None else None
        d_xyz, d_rot, _ = sw.skin(self.pc.get_xyz, joints, self.warp.parents_i32, self.warp._node_radius, G, node_rot,
                                  pose["global_trans"], mask, self.mlp_weight, self.offsets, 0, False)
        if self.isotropic:  # the trainer zeroes d_rotation for isotropic Gaussians (train_rig.py:422-423)
            d_rot = 0.0
        return render(cam, self.pc, self.pipe, self.bg, d_xyz, d_rot, 0.0)

    def step(self, cam, pose, target: torch.Tensor) -> torch.Tensor:
        """Forward + backward of one view; returns the (device) loss.  Gradients accumulate in .grad."""
        out = self.forward(cam, pose)
        loss = (out["render"] - target).abs().mean()
        loss.backward()
        return loss.detach()
