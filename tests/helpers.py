"""Shared helpers of the test-suite: scene -> rasterizer inputs, and runners for the three implementations
(ours through the C ABI, the CPU oracle, the compiled reference CUDA)."""
from __future__ import annotations

import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import oracle_py as orc  # noqa: E402
from oracle import skin_oracle as so  # noqa: E402
from riggs_b200 import synth  # noqa: E402


def raster_inputs(scene, view=0, pose=None, bg=(0.0, 0.0, 0.0), skinned=True):
    """Activated rasterizer inputs (CPU tensors) for one view of a scene, deformed by pose `pose`."""
    cam = scene.cameras[view]
    if skinned:
        p = scene.poses[view if pose is None else pose]
        d = so.deform_by_pose(scene.xyz, scene.joints, scene.parents, scene.node_radius_raw, p["local_rotation"],
                              p["global_trans"], scene.motion_mask, scene.mlp_weight, scene.template_offsets)
        d_xyz, d_rot = d["d_xyz"], d["d_rotation"]
    else:
        d_xyz, d_rot = 0.0, 0.0
    means3D, scales, rots, opac = so.activate(scene.xyz, scene.scaling_raw, scene.rotation_raw, scene.opacity_raw,
                                              d_xyz, d_rot)
    shs = torch.cat([scene.features_dc, scene.features_rest], 1)
    return dict(means3D=means3D.contiguous(), scales=scales.contiguous(), rotations=rots.contiguous(),
                opacities=opac.contiguous(), shs=shs.contiguous(), view=cam.world_view_transform.contiguous(),
                proj=cam.full_proj_transform.contiguous(), campos=cam.camera_center.contiguous(),
                W=cam.image_width, H=cam.image_height, tanfovx=math.tan(cam.FoVx * 0.5),
                tanfovy=math.tan(cam.FoVy * 0.5), bg=torch.tensor(bg, dtype=torch.float32))


def np_in(inp):
    return {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in inp.items()}


def oracle_forward(inp, **kw):
    a = np_in(inp)
    return orc.forward(a["means3D"], a["opacities"], a["view"], a["proj"], a["campos"], a["W"], a["H"], a["tanfovx"],
                       a["tanfovy"], a["bg"], shs=a.get("shs"), colors_precomp=a.get("colors_precomp"),
                       scales=a.get("scales"), rotations=a.get("rotations"), cov3D_precomp=a.get("cov3D_precomp"),
                       degree=a.get("degree", 3), **kw)


def oracle_backward(st, inp, cots):
    a = np_in(inp)
    gc, gd, ga = (c.numpy() if torch.is_tensor(c) else c for c in cots)
    return orc.backward(st, a["means3D"], a["view"], a["proj"], a["campos"], a["W"], a["H"], a["tanfovx"],
                        a["tanfovy"], a["bg"], gc, gd, ga, shs=a.get("shs"), colors_precomp=a.get("colors_precomp"),
                        scales=a.get("scales"), rotations=a.get("rotations"), cov3D_precomp=a.get("cov3D_precomp"),
                        degree=a.get("degree", 3))


def to_cuda(inp):
    return {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in inp.items()}


def _slice(buf, addr, nbytes, dtype):
    off = addr - buf.data_ptr()
    return buf[off:off + nbytes].view(dtype)


def ours_forward(cin):
    """Run libriggs_b200 through the reference-shaped binding; returns outputs + decoded state (numpy)."""
    from riggs_b200.diff_gaussian_rasterization import _C
    e = torch.empty(0)
    g = lambda k: cin[k] if cin.get(k) is not None else e
    R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(
        cin["bg"], cin["means3D"], g("colors_precomp"), cin["opacities"], g("scales"), g("rotations"), 1.0,
        g("cov3D_precomp"), cin["view"], cin["proj"], cin["tanfovx"], cin["tanfovy"], cin["H"], cin["W"], g("shs"),
        cin.get("degree", 3), cin["campos"], False, False)
    P, W, H = cin["means3D"].shape[0], cin["W"], cin["H"]
    T = ((W + 15) // 16) * ((H + 15) // 16)
    ptrs = _C.state_ptrs(P, R, W, H, geom, binning if R > 0 else None, img)
    st = dict(num_rendered=R, color=color, depth=depth, alpha=alpha, radii_t=radii, geom=geom, binning=binning, img=img)
    st["radii"] = radii.cpu().numpy()
    st["depths"] = _slice(geom, ptrs[0], 4 * P, torch.float32).cpu().numpy()
    st["clamped_bits"] = _slice(geom, ptrs[1], P, torch.uint8).cpu().numpy()
    st["means2D"] = _slice(geom, ptrs[2], 8 * P, torch.float32).cpu().numpy().reshape(P, 2)
    st["conic_opacity"] = _slice(geom, ptrs[3], 16 * P, torch.float32).cpu().numpy().reshape(P, 4)
    st["rgb"] = _slice(geom, ptrs[4], 64 * P, torch.float32).cpu().numpy().reshape(P, 16)[:, 8:11].copy()
    st["tiles_touched"] = _slice(geom, ptrs[5], 4 * P, torch.int32).cpu().numpy().view(np.uint32)
    st["sorted_ids"] = _slice(geom, ptrs[6], 4 * P, torch.int32).cpu().numpy().view(np.uint32)
    if R > 0:
        st["point_list"] = _slice(binning, ptrs[8], 4 * R, torch.int32).cpu().numpy().view(np.uint32)
        st["keys"] = _C.debug_sorted_keys(P, R, W, H, geom, binning, img).cpu().numpy().view(np.uint64)
    else:
        st["point_list"] = np.zeros(0, np.uint32)
        st["keys"] = np.zeros(0, np.uint64)
    st["n_contrib"] = _slice(img, ptrs[9], 4 * W * H, torch.int32).cpu().numpy().view(np.uint32).reshape(H, W)
    st["ranges"] = _slice(img, ptrs[10], 8 * T, torch.int32).cpu().numpy().view(np.uint32).reshape(T, 2)
    return st


def ours_backward(cin, st, cots):
    from riggs_b200.diff_gaussian_rasterization import _C
    e = torch.empty(0)
    g = lambda k: cin[k] if cin.get(k) is not None else e
    gc, gd, ga = (c.cuda() for c in cots)
    out = _C.rasterize_gaussians_backward(
        cin["bg"], cin["means3D"], st["radii_t"], g("colors_precomp"), g("scales"), g("rotations"), 1.0,
        g("cov3D_precomp"), cin["view"], cin["proj"], cin["tanfovx"], cin["tanfovy"], gc, gd, ga, g("shs"),
        cin.get("degree", 3), cin["campos"], st["geom"], st["num_rendered"], st["binning"], st["img"], st["alpha"],
        False)
    names = ["dmean2D", "dcolor", "dopacity", "dmean3D", "dcov3D", "dsh", "dscale", "drot"]
    return dict(zip(names, out))


def rel_err(a, b, floor=0.0):
    """max |a-b| / max(|b|_inf, floor): error relative to the scale of the reference array."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), floor, 1e-30))


def rel_l2(a, b):
    """|a - b|_2 / |b|_2: pins the bulk of the entries, which a max-norm ratio (rel_err) lets through."""
    a = np.asarray(a, np.float64).reshape(-1)
    b = np.asarray(b, np.float64).reshape(-1)
    if a.size == 0:
        return 0.0
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def floored_rel(a, b, floor_frac=1e-3):
    """max over the entries of |a - b| / max(|b|, floor_frac * max|b|): every entry down to floor_frac of the largest
    one is compared on its OWN scale, smaller ones on the floor's."""
    a = np.asarray(a, np.float64).reshape(-1)
    b = np.asarray(b, np.float64).reshape(-1)
    if a.size == 0:
        return 0.0
    floor = max(float(np.abs(b).max()) * floor_frac, 1e-30)
    return float((np.abs(a - b) / np.maximum(np.abs(b), floor)).max())


def grad_stats(a, b):
    return dict(max=rel_err(a, b), l2=rel_l2(a, b), floored=floored_rel(a, b))
