"""skin_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU (torch) restatement of the reference's skeleton skinning path, in the reference's own
eager-op style so that, timed on the host cores, it is a fair stand-in for the reference's
torch CPU skinning (the reference tree itself cannot travel to the GPU box).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.

Pinned against the real reference: tests/golden/make_skin_golden.py imports
``skeleton_utils.skeleton_warp.SkeletonWarp`` and ``utils.dual_quaternion`` from
/root/reference on the CPU and stores inputs + outputs (and autograd gradients) in
tests/golden/skin_*.npz; tests/test_oracle_skin.py replays them through this file.

Reference lines restated:
  quaternion_to_matrix          utils/time_utils.py:115-132
  matrix_to_quaternion          utils/time_utils.py:146-205
  line_segment_distance         skeleton_utils/skeleton_warp.py:215-238
  cal_nn_weight_skeleton (K=-1) skeleton_utils/skeleton_warp.py:41-76
  chain_product_transform       skeleton_utils/skeleton_warp.py:242-273
  deform_by_pose                skeleton_utils/skeleton_warp.py:130-172
  QT2DQ / DQ2QT / DQBlending    utils/dual_quaternion.py:135-180
  render() activations          gaussian_renderer/__init__.py:74-92, scene/gaussian_model.py:105-132
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def quat_to_mat(q: torch.Tensor) -> torch.Tensor:
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def mat_to_quat(m: torch.Tensor) -> torch.Tensor:
    """Four candidates, pick the one with the largest |component|; no sign standardisation."""
    b = m.shape[:-2]
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(m.reshape(b + (9,)), -1)
    t = torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22, 1.0 - m00 + m11 - m22,
                     1.0 - m00 - m11 + m22], -1)
    q_abs = torch.where(t > 0, torch.sqrt(t.clamp_min(0)), torch.zeros_like(t))
    cand = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], -1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], -1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], -1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], -1)], -2)
    cand = cand / (2.0 * q_abs[..., None].clamp_min(0.1))
    pick = q_abs.argmax(-1)
    return torch.gather(cand, -2, pick[..., None, None].expand(b + (1, 4))).squeeze(-2)


def bone_dist2(x: torch.Tensor, joints: torch.Tensor, parents: torch.Tensor) -> torch.Tensor:
    """(N, J-1) squared distance from each point to bone segment parent(k) -> k, k = 1..J-1."""
    b = joints[1:, :3]
    a = joints[parents[1:].long(), :3]
    p = x[:, None, :3]
    ab = b - a
    t = ((p - a) * ab).sum(-1, keepdim=True) / torch.clamp((ab * ab).sum(-1, keepdim=True), min=1e-6)
    t = torch.clamp(t, 0.0, 1.0)
    s = a + t * ab
    return ((s - p) * (s - p)).sum(-1)


def skin_weights(x, joints, parents, node_radius_raw, mlp_weight=None, K=-1, nn_idx=None):
    """cal_nn_weight_skeleton (skeleton_warp.py:41-76).  K <= 0: all bones, returns (w (N,J-1), d2 (N,J-1), None);
    K > 0: the K nearest bones (torch.topk, ascending), returns (w (N,K), d2 (N,K), idx (N,K) = joint numbers 1..J-1)."""
    d2 = bone_dist2(x, joints, parents)
    idx = None
    if K > 0:
        if nn_idx is not None:  # selection given (tests: equidistant bones make torch.topk's choice arbitrary)
            idx = nn_idx.long() - 1
            d2 = torch.gather(d2, 1, idx)
        else:
            d2, idx = torch.topk(d2, K, largest=False, dim=1)
        idx = idx + 1
        radius = torch.exp(node_radius_raw)[idx]
        if mlp_weight is not None:
            raise ValueError("K > 0 with the weight MLP is out of range in the reference (skeleton_warp.py:59)")
    else:
        radius = torch.exp(node_radius_raw)[1:]
    w = torch.exp(-d2 / (2 * radius ** 2))
    if mlp_weight is not None:
        w = w * mlp_weight
    w = w + 1e-7
    return w / w.sum(-1, keepdim=True), d2, idx


def forward_kinematics(local_rot_mats: torch.Tensor, joints: torch.Tensor, parents: torch.Tensor):
    """Returns posed joints (J,3) and global transforms (J,4,4)."""
    J = joints.shape[0]
    vp = parents.clone().long()
    vp[0] = 0
    pj = joints[vp]
    local_t = pj - torch.einsum("jab,jb->ja", local_rot_mats, pj)
    T = torch.zeros(J, 4, 4, dtype=joints.dtype, device=joints.device)
    T[:, :3, :3] = local_rot_mats
    T[:, :3, 3] = local_t
    T[:, 3, 3] = 1.0
    chain = [T[0]]
    for i in range(1, J):
        chain.append(chain[int(parents[i])] @ T[i])
    G = torch.stack(chain, 0)
    posed = torch.einsum("jab,jb->ja", G[:, :3, :3], joints) + G[:, :3, 3]
    return posed, G


def deform_by_pose(x, joints, parents, node_radius_raw, local_rotation, global_trans, motion_mask,
                   mlp_weight=None, template_offsets=None, K=-1, nn_idx=None, blend_mode="lbs"):
    """LBS core.  Returns the dict of the reference's third boundary.
    blend_mode "dq": the bone transforms (node_rot as detached quaternions, the translations of G) are blended as dual
    quaternions (utils/dual_quaternion.py:135-180, DQBlending) instead of linearly; d_rotation is then the blended unit
    quaternion times the mask.  K = -1 only."""
    x = x.detach()
    R_local = quat_to_mat(local_rotation)
    w, d2, idx = skin_weights(x, joints, parents, node_radius_raw, mlp_weight, K, nn_idx)
    posed, G = forward_kinematics(R_local, joints[:, :3], parents)
    G_R, G_t = G[:, :3, :3], G[:, :3, 3]
    node_rot = mat_to_quat(G_R.detach())
    if idx is None:
        Ax = torch.einsum("kab,nb->nka", G_R[1:], x) + G_t[1:][None]
        rot = node_rot[1:][None]
        nn_idx = torch.arange(1, joints.shape[0], device=x.device).unsqueeze(0).expand(x.shape[0], -1)
    else:
        Ax = torch.einsum("nkab,nb->nka", G_R[idx], x) + G_t[idx]
        rot = node_rot[idx]
        nn_idx = idx
    Ax_avg = (Ax * w[..., None]).sum(1)
    dq_rot = None
    if blend_mode == "dq":
        if idx is not None:
            raise ValueError("dual-quaternion blending is defined for K = -1 (all bones)")
        Rb, Tb = dq_blend(node_rot[1:], G_t[1:], w)
        Ax_avg = torch.einsum("nab,nb->na", Rb, x) + Tb
        dq_rot = dq_blend_quat(node_rot[1:], G_t[1:], w)
    Ax_avg = Ax_avg + global_trans
    if template_offsets is not None:
        Ax_avg = Ax_avg + template_offsets
    d_xyz = (Ax_avg - x) * motion_mask
    d_rot = (rot * w[..., None]).sum(1) * motion_mask
    if dq_rot is not None:
        d_rot = dq_rot * motion_mask
    return {"d_xyz": d_xyz, "d_rotation": d_rot,
            "d_scaling": torch.zeros(x.shape[0], 3, dtype=x.dtype, device=x.device),
            "d_nodes": posed + global_trans, "nn_weight": w, "nn_idx": nn_idx, "nn_dist2": d2,
            "G": G, "node_rot": node_rot}


# ---- dual-quaternion blending (utils/dual_quaternion.py) -------------------------------------

def _qmul_raw(a, b):
    aw, ax, ay, az = torch.unbind(a, -1)
    bw, bx, by, bz = torch.unbind(b, -1)
    return torch.stack((aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw), -1)


def _qmul(a, b):
    ab = _qmul_raw(a, b)
    return torch.where(ab[..., 0:1] < 0, -ab, ab)


def qt_to_dq(q, t):
    q = F.normalize(q)  # note: dim=1, as the reference calls it (dual_quaternion.py:138)
    tq = torch.cat([torch.zeros_like(t[..., :1]), t], -1)
    return torch.cat([q, _qmul(tq, q) / 2], -1)


def dq_to_rt(dq):
    real, imag = dq[..., :4], dq[..., 4:]
    n = real.norm(dim=-1, keepdim=True).clamp(min=1e-8)
    real, imag = real / n, imag / n
    w0, x0, y0, z0 = torch.unbind(real, -1)
    w1, x1, y1, z1 = torch.unbind(imag, -1)
    t = 2 * torch.stack([-w1 * x0 + x1 * w0 - y1 * z0 + z1 * y0, -w1 * y0 + x1 * z0 + y1 * w0 - z1 * x0,
                         -w1 * z0 - x1 * y0 + y1 * x0 + z1 * w0], -1)
    R = torch.stack([1 - 2 * y0 ** 2 - 2 * z0 ** 2, 2 * x0 * y0 - 2 * w0 * z0, 2 * x0 * z0 + 2 * w0 * y0,
                     2 * x0 * y0 + 2 * w0 * z0, 1 - 2 * x0 ** 2 - 2 * z0 ** 2, 2 * y0 * z0 - 2 * w0 * x0,
                     2 * x0 * z0 - 2 * w0 * y0, 2 * y0 * z0 + 2 * w0 * x0, 1 - 2 * x0 ** 2 - 2 * y0 ** 2],
                    -1).reshape(*w0.shape, 3, 3)
    return R, t


def dq_blend(q, t, weights):
    """q (K,4) already-normalised per-bone quats, t (K,3), weights (N,K) -> R (N,3,3), t (N,3)."""
    dq = qt_to_dq(q, t)
    avg = (dq[None] * weights[..., None]).sum(-2)
    return dq_to_rt(avg)


def dq_blend_quat(q, t, weights):
    """The blended unit quaternion itself (real part of the normalised blend; DQ2QT's `real`, dual_quaternion.py:148-151)."""
    dq = qt_to_dq(q, t)
    avg = (dq[None] * weights[..., None]).sum(-2)
    real = avg[..., :4]
    return real / real.norm(dim=-1, keepdim=True).clamp(min=1e-8)


# ---- render() activations --------------------------------------------------------------------

def activate(xyz, scaling_raw, rotation_raw, opacity_raw, d_xyz, d_rotation, d_scaling=0.0):
    means3D = xyz + d_xyz
    s = scaling_raw if scaling_raw.shape[-1] == 3 else scaling_raw[..., :1].repeat(1, 3)
    scales = torch.exp(s) + d_scaling
    rotations = F.normalize(rotation_raw + d_rotation)
    opacity = torch.sigmoid(opacity_raw)
    return means3D, scales, rotations, opacity


# ---- stage-1 control-node warp ------------------------------------------------------
# Reference lines restated: ControlNodeWarp.cal_nn_weight utils/time_utils.py:934-967 (non-skinning branch,
# gs_kernel=True) and ControlNodeWarp.forward :1133-1236 without the animation branches (node_trans_bias,
# animation_d_values) and the ARAP regulariser.  Pinned by tests/golden/node_warp_*.npz, which
# tests/golden/make_node_warp_golden.py produced from the reference's own methods.

def knn_brute(p, nodes, K):
    """pytorch3d.ops.knn_points' contract for one cloud: K smallest squared distances, ascending, differentiable."""
    d = ((p[:, None, :] - nodes[None, :, :]) ** 2).sum(-1)
    idx = torch.topk(d.detach().double(), K, dim=-1, largest=False, sorted=True).indices
    return torch.gather(d, 1, idx), idx


def node_nn_weight(x, feature, nodes, node_radius_raw, node_weight_raw, K, hyper_dim):
    """(nn_weight, nn_dist, nn_idx): utils/time_utils.py:943-963."""
    p = x.detach()
    if hyper_dim > 0 and feature is not None:
        p = torch.cat([p, feature[..., :hyper_dim]], -1)
        nd = torch.cat([nodes[..., :3].detach(), nodes[..., 3:]], -1)
    else:
        nd = nodes[..., :3].detach()
    nn_dist, nn_idx = knn_brute(p, nd, K)
    radius = torch.exp(node_radius_raw)[nn_idx]
    w = torch.exp(-nn_dist / (2 * radius ** 2))
    if node_weight_raw is not None:
        w = w * torch.sigmoid(node_weight_raw.reshape(-1))[nn_idx]
    w = w + 1e-7
    return w / w.sum(-1, keepdim=True), nn_dist, nn_idx


def node_warp(x, feature, nodes, node_radius_raw, node_weight_raw, node_attrs, motion_mask, K=3, hyper_dim=2,
              local_frame=True, d_rot_as_res=True):
    """ControlNodeWarp.forward (:1133-1198, 1213-1224): blends the per-node motion onto the points."""
    x = x.detach()
    w, dist, idx = node_nn_weight(x, feature, nodes, node_radius_raw, node_weight_raw, K, hyper_dim)
    bias = torch.tensor([1.0, 0, 0, 0], dtype=x.dtype)
    t, r, s = node_attrs["d_xyz"], node_attrs["d_rotation"], node_attrs["d_scaling"]
    if local_frame:
        R = quat_to_mat(node_attrs["local_rotation"] + bias)
        c = nodes[idx][..., :3].detach()
        Ax = torch.einsum("nkab,nkb->nka", R[idx], x[:, None] - c) + c + t[idx]
        translate = (Ax * w[..., None]).sum(1) - x
    else:
        translate = (t[idx] * w[..., None]).sum(1)
    translate = translate * motion_mask
    if d_rot_as_res:
        rotation = (r[idx] * w[..., None]).sum(1) * motion_mask
    else:
        rotation = (((r + bias)[idx] * w[..., None]).sum(1) - bias) * motion_mask + bias
    out = {"d_xyz": translate, "d_rotation": rotation, "d_scaling": (s[idx] * w[..., None]).sum(1) * motion_mask,
           "d_nodes": nodes[..., :3] + t, "nn_weight": w, "nn_dist": dist, "nn_idx": idx,
           "d_opacity": None, "d_color": None}
    for key in ("d_opacity", "d_color"):
        if node_attrs.get(key) is not None:
            out[key] = (node_attrs[key][idx] * w[..., None]).sum(1) * motion_mask
    return out
