"""``render()`` with the reference's signature (gaussian_renderer/__init__.py:37-151), on the B200 kernels.

``pc`` is anything with the accessors the reference's GaussianModel offers (scene/gaussian_model.py:95-142):
``get_xyz, get_opacity, get_scaling, get_rotation_bias(d_rot), get_features, active_sh_degree, max_sh_degree,
motion_mask, get_covariance`` -- the reference's own GaussianModel works unchanged, and
``riggs_b200.gaussian_renderer.GaussianParams`` is a minimal holder with the same accessors for tests and
benchmarks.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

C0 = 0.28209479177387814


def standardize_quaternion(q):
    return torch.where(q[..., 0:1] < 0, -q, q)


def quaternion_raw_multiply(a, b):
    aw, ax, ay, az = torch.unbind(a, -1)
    bw, bx, by, bz = torch.unbind(b, -1)
    return torch.stack((aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw), -1)


def quaternion_multiply(a, b):
    return standardize_quaternion(quaternion_raw_multiply(a, b))


class GaussianParams(torch.nn.Module):
    """Raw Gaussian parameters + the activations of scene/gaussian_model.py:95-132."""

    def __init__(self, xyz, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest, motion_mask=None,
                 sh_degree=3, device="cuda"):
        super().__init__()
        P = torch.nn.Parameter
        self._xyz = P(xyz.float().to(device))
        self._scaling = P(scaling_raw.float().to(device))
        self._rotation = P(rotation_raw.float().to(device))
        self._opacity = P(opacity_raw.float().to(device))
        self._features_dc = P(features_dc.float().to(device))
        self._features_rest = P(features_rest.float().to(device))
        self._motion_mask = motion_mask.float().to(device) if motion_mask is not None else None
        self.active_sh_degree = sh_degree
        self.max_sh_degree = 3
        self.use_isotropic_gs = scaling_raw.shape[-1] == 1

    @classmethod
    def from_scene(cls, scene, device="cuda"):
        return cls(scene.xyz, scene.scaling_raw, scene.rotation_raw, scene.opacity_raw, scene.features_dc,
                   scene.features_rest, scene.motion_mask, scene.sh_degree, device)

    @property
    def motion_mask(self):
        return self._motion_mask if self._motion_mask is not None else torch.ones_like(self._xyz[..., :1])

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_scaling(self):
        if self.use_isotropic_gs:
            return torch.exp(self._scaling[..., :1].repeat(1, 3))
        return torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return F.normalize(self._rotation)

    def get_rotation_bias(self, rotation_bias=None):
        rotation_bias = rotation_bias if rotation_bias is not None else 0.
        return F.normalize(self._rotation + rotation_bias)

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_opacity(self):
        return torch.sigmoid(self._opacity)


def eval_sh_python(deg, sh, dirs):
    """utils/sh_utils.py:57-112 for the convert_SHs_python flag; sh (...,C,(deg+1)^2)."""
    C1 = 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435]
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] +
                      C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7] + C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                          C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                          C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                          C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return result


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, d_xyz, d_rotation, d_scaling, d_opacity=None,
           d_color=None, scaling_modifier=1.0, override_color=None, random_bg_color=False, render_motion=False,
           detach_xyz=False, detach_scale=False, detach_rot=False, detach_opacity=False, d_rot_as_res=True,
           scale_const=None, d_rotation_bias=None, force_visible=False):
    """Render the scene; the background tensor must be on the GPU.  Same dict as the reference (:145-151)."""
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    bg = bg_color if not random_bg_color else torch.rand_like(bg_color)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False,
        debug=getattr(pipe, "debug", False))
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    means3D = pc.get_xyz + d_xyz
    means2D = screenspace_points
    if scale_const is not None:
        opacity = torch.ones_like(pc.get_opacity)
    else:
        opacity = pc.get_opacity if d_opacity is None else pc.get_opacity + d_opacity

    scales = rotations = cov3D_precomp = None
    if getattr(pipe, "compute_cov3D_python", False):
        cov3D_precomp = pc.get_covariance(scaling_modifier, d_rotation=None if type(d_rotation) is float else d_rotation,
                                          gs_rot_bias=d_rotation_bias)
    else:
        scales = pc.get_scaling + d_scaling
        rotations = pc.get_rotation_bias(d_rotation)
        if d_rotation_bias is not None:
            rotations = quaternion_multiply(d_rotation_bias, rotations)

    shs = colors_precomp = None
    if render_motion:
        colors_precomp = torch.zeros_like(pc.get_xyz)
        colors_precomp[..., :1] = pc.motion_mask
        colors_precomp[..., -1:] = 1 - pc.motion_mask
    elif override_color is None:
        if d_color is not None and type(d_color) is not float:
            sh_features = torch.cat([pc.get_features[:, :1] + d_color[:, None], pc.get_features[:, 1:]], dim=1)
        else:
            sh_features = pc.get_features
        if getattr(pipe, "convert_SHs_python", False):
            shs_view = sh_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = pc.get_xyz - viewpoint_camera.camera_center.repeat(sh_features.shape[0], 1)
            dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            sh2rgb = eval_sh_python(pc.active_sh_degree, shs_view, dir_pp_normalized)
            colors_precomp = torch.clamp_min(sh2rgb + 0.5, 0.0)
        else:
            shs = sh_features
    else:
        colors_precomp = override_color

    if detach_xyz:
        means3D = means3D.detach()
    if detach_rot or detach_scale:
        if cov3D_precomp is not None:
            cov3D_precomp = cov3D_precomp.detach()
        else:
            rotations = rotations.detach() if detach_rot else rotations
            scales = scales.detach() if detach_scale else scales
    if detach_opacity:
        opacity = opacity.detach()
    if scale_const is not None:
        scales = scale_const * torch.ones_like(scales)

    rendered_image, radii, depth, alpha = rasterizer(means3D=means3D, means2D=means2D, shs=shs,
                                                     colors_precomp=colors_precomp, opacities=opacity, scales=scales,
                                                     rotations=rotations, cov3D_precomp=cov3D_precomp)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "depth": depth, "alpha": alpha, "bg_color": bg}


def render_flow(pc, viewpoint_camera1, viewpoint_camera2, d_xyz1, d_xyz2, d_rotation1, d_scaling1, scaling_modifier=1.0,
                compute_cov3D_python=False, scale_const=None, d_rot_as_res=True, **kwargs):
    """Splat the per-Gaussian screen-space flow between two deformations (and the motion mask in the third channel) with
    the Gaussians posed by the first one.  Same signature, arithmetic and result dict as the reference's ``render_flow``
    (gaussian_renderer/__init__.py:154-269): the flow is ``project(xyz + d_xyz2, cam2) - project(xyz + d_xyz1, cam1)``
    on detached positions, rasterised as ``colors_precomp`` over a zero background with SH degree 0."""
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera1.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera1.FoVy * 0.5)

    canonical = pc.get_xyz.clone().detach()  # the flow itself carries no gradient to the positions (:187)
    xyz_t1 = canonical + d_xyz1
    xyz_t2 = canonical + d_xyz2

    def project(xyz, cam):
        hom = torch.cat([xyz, torch.ones_like(xyz[..., :1])], dim=-1) @ cam.full_proj_transform
        return hom[..., :3] / hom[..., -1:]

    cam2 = viewpoint_camera2 if viewpoint_camera2 is not None else viewpoint_camera1
    flow = project(xyz_t2, cam2) - project(xyz_t1, viewpoint_camera1)
    flow[..., -1:] = pc.motion_mask  # third channel: the motion mask (:203)

    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera1.image_height), image_width=int(viewpoint_camera1.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=torch.zeros_like(flow[0]), scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera1.world_view_transform, projmatrix=viewpoint_camera1.full_proj_transform,
        sh_degree=0, campos=viewpoint_camera1.camera_center, prefiltered=False, debug=False)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    means3D = pc.get_xyz + d_xyz1
    opacity = pc.get_opacity

    def posed_rotation():
        if d_rot_as_res:
            return pc.get_rotation + d_rotation1
        return pc.get_rotation if type(d_rotation1) is float else quaternion_multiply(d_rotation1, pc.get_rotation)

    scales = rotations = cov3D_precomp = None
    if scale_const is not None:
        scales = torch.ones_like(pc.get_scaling) * scale_const
        rotations = posed_rotation()
    elif compute_cov3D_python:
        cov3D_precomp = pc.get_covariance(scaling_modifier, d_rotation=None if type(d_rotation1) is float else d_rotation1)
    else:
        scales = pc.get_scaling + d_scaling1
        rotations = posed_rotation()

    rendered_image, radii, depth, alpha = rasterizer(means3D=means3D, means2D=screenspace_points, shs=None,
                                                     colors_precomp=flow, opacities=opacity, scales=scales,
                                                     rotations=rotations, cov3D_precomp=cov3D_precomp)
    return {"render": rendered_image, "depth": depth, "alpha": alpha, "viewspace_points": screenspace_points,
            "visibility_filter": radii > 0, "radii": radii}
