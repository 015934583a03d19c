"""Seeded synthetic scenes for the skin + splat-rasterize hot path.

No dataset or checkpoint is reachable from the build machines, so every test and
every bench line runs on scenes produced here.  The shapes follow the things the
reference feeds its hot path:

* skeleton tree in BFS order (``parents[0] == -1``, ``parents[i] < i``) as produced
  by ``skeleton_utils/extract_skeleton_utils.py:32-57``;
* Gaussians scattered around the bones, raw (pre-activation) parameters laid out
  like ``scene/gaussian_model.py:105-132`` expects them;
* one pose per frame in the form ``SkeletonWarp.forward`` hands to
  ``deform_by_pose`` (``skeleton_utils/skeleton_warp.py:119-121``):
  ``local_rotation = [1,0,0,0] + small``, ``global_trans`` small;
* cameras with the matrix conventions of ``scene/cameras.py:55-72`` and
  ``utils/graphics_utils.py:42-76`` (both matrices stored transposed, i.e. the
  row-vector convention the rasterizer reads as ``m[4*c + r]``).

Everything is generated on the CPU with a fixed ``torch.Generator`` so the same
config id gives the same bytes on every machine.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

SMPL_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]


@dataclass
class Camera:
    """The fields of ``scene.cameras.Camera`` that ``render()`` reads."""

    image_width: int
    image_height: int
    FoVx: float
    FoVy: float
    world_view_transform: torch.Tensor  # (4,4), transposed storage
    full_proj_transform: torch.Tensor  # (4,4), transposed storage
    camera_center: torch.Tensor  # (3,)

    def to(self, device):
        return Camera(self.image_width, self.image_height, self.FoVx, self.FoVy,
                      self.world_view_transform.to(device), self.full_proj_transform.to(device),
                      self.camera_center.to(device))


@dataclass
class Scene:
    """Raw parameters of one synthetic rigged-Gaussian scene (all fp32, CPU)."""

    joints: torch.Tensor  # (J,3) rest joints
    parents: torch.Tensor  # (J,) int64
    node_radius_raw: torch.Tensor  # (J,) log radius (``_node_radius``)
    xyz: torch.Tensor  # (N,3)
    scaling_raw: torch.Tensor  # (N,3) or (N,1): log scale
    rotation_raw: torch.Tensor  # (N,4)
    opacity_raw: torch.Tensor  # (N,1) logit
    features_dc: torch.Tensor  # (N,1,3)
    features_rest: torch.Tensor  # (N,15,3)
    motion_mask: torch.Tensor  # (N,1)
    sh_degree: int = 3
    cameras: List[Camera] = field(default_factory=list)
    poses: List[Dict[str, torch.Tensor]] = field(default_factory=list)
    mlp_weight: Optional[torch.Tensor] = None  # (N,J-1) sigmoid(WeightMLP(x)) stand-in
    template_offsets: Optional[torch.Tensor] = None  # (N,3) detail_net stand-in

    @property
    def N(self) -> int:
        return self.xyz.shape[0]

    @property
    def J(self) -> int:
        return self.joints.shape[0]


def _gen(seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return g


def make_skeleton(J: int, g: torch.Generator):
    """Rest skeleton.  J == 24 gives an SMPL-like figure about 1.7 units tall."""
    if J == 24:
        parents = torch.tensor(SMPL_PARENTS, dtype=torch.int64)
        rest = torch.tensor([
            [0.00, 0.00, 0.95], [0.09, 0.00, 0.86], [-0.09, 0.00, 0.86], [0.00, 0.00, 1.07],
            [0.10, 0.00, 0.48], [-0.10, 0.00, 0.48], [0.00, 0.00, 1.20], [0.10, 0.00, 0.08],
            [-0.10, 0.00, 0.08], [0.00, 0.00, 1.27], [0.11, 0.12, 0.02], [-0.11, 0.12, 0.02],
            [0.00, 0.00, 1.48], [0.08, 0.00, 1.40], [-0.08, 0.00, 1.40], [0.00, 0.03, 1.60],
            [0.19, 0.00, 1.42], [-0.19, 0.00, 1.42], [0.45, 0.00, 1.42], [-0.45, 0.00, 1.42],
            [0.70, 0.00, 1.42], [-0.70, 0.00, 1.42], [0.79, 0.00, 1.42], [-0.79, 0.00, 1.42],
        ], dtype=torch.float32)
        rest[:, 2] -= 0.85  # centre the figure on the origin
        return rest, parents
    parents = torch.full((J,), -1, dtype=torch.int64)
    joints = torch.zeros(J, 3)
    for i in range(1, J):
        lo = max(0, i - 4)
        p = int(torch.randint(lo, i, (1,), generator=g))
        parents[i] = p
        d = torch.randn(3, generator=g)
        d = d / d.norm().clamp_min(1e-6)
        length = 0.15 + 0.20 * float(torch.rand(1, generator=g))
        joints[i] = joints[p] + d * length
    joints -= joints.mean(0, keepdim=True)
    return joints, parents


def look_at_camera(eye: torch.Tensor, width: int, height: int, fovx: float,
                   znear: float = 0.01, zfar: float = 100.0) -> Camera:
    """Camera at ``eye`` looking at the origin, +z forward, matrices stored transposed."""
    eye = eye.double()
    fwd = -eye / eye.norm()
    up = torch.tensor([0.0, 0.0, 1.0], dtype=torch.float64)
    if abs(float(fwd @ up)) > 0.99:
        up = torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64)
    right = torch.linalg.cross(fwd, up)
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    R_w2c = torch.stack([right, down, fwd], 0)  # rows are the camera axes in world coords
    W2C = torch.eye(4, dtype=torch.float64)
    W2C[:3, :3] = R_w2c
    W2C[:3, 3] = -R_w2c @ eye
    fovy = 2.0 * math.atan(math.tan(fovx * 0.5) * height / width)
    tx, ty = math.tan(fovx * 0.5), math.tan(fovy * 0.5)
    Pm = torch.zeros(4, 4, dtype=torch.float64)
    Pm[0, 0] = 1.0 / tx
    Pm[1, 1] = 1.0 / ty
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    wv = W2C.float().t().contiguous()
    proj = Pm.float().t().contiguous()
    full = (wv.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wv.inverse()[3, :3].contiguous()
    return Camera(width, height, fovx, fovy, wv, full, center)


def make_cameras(n_views: int, width: int, height: int, g: torch.Generator,
                 radius: float = 4.0, fovx: float = 0.6911) -> List[Camera]:
    cams = []
    jitter = float(torch.rand(1, generator=g)) * 2 * math.pi
    for k in range(n_views):
        az = 2 * math.pi * k / max(n_views, 1) + jitter
        el = math.radians(-30.0 + 60.0 * float(torch.rand(1, generator=g)))
        eye = torch.tensor([radius * math.cos(el) * math.cos(az), radius * math.cos(el) * math.sin(az),
                            radius * math.sin(el)])
        cams.append(look_at_camera(eye, width, height, fovx))
    return cams


def make_pose(J: int, g: torch.Generator, rot_sigma: float = 0.15, trans_sigma: float = 0.05):
    q = torch.zeros(J, 4)
    q[:, 0] = 1.0
    q = q + rot_sigma * torch.randn(J, 4, generator=g)
    t = trans_sigma * torch.randn(3, generator=g)
    return {"local_rotation": q, "global_trans": t}


def slerp_pose_sequence(J: int, n_frames: int, g: torch.Generator, n_keys: int = 9):
    """Novel-pose sequence: spherical interpolation between key poses
    (the shape of ``skeleton_utils/interpolation_utils.py:58-86``)."""
    keys = [make_pose(J, g, rot_sigma=0.25) for _ in range(n_keys)]
    per = max(1, n_frames // (n_keys - 1))
    out = []
    for f in range(n_frames):
        k = min(f // per, n_keys - 2)
        u = (f - k * per) / per
        q0 = torch.nn.functional.normalize(keys[k]["local_rotation"], dim=-1)
        q1 = torch.nn.functional.normalize(keys[k + 1]["local_rotation"], dim=-1)
        dot = (q0 * q1).sum(-1, keepdim=True)
        q1 = torch.where(dot < 0, -q1, q1)
        dot = dot.abs().clamp(max=1.0)
        th = torch.acos(dot)
        s = torch.sin(th)
        w0 = torch.where(s > 1e-4, torch.sin((1 - u) * th) / s.clamp_min(1e-8), torch.full_like(s, 1 - u))
        w1 = torch.where(s > 1e-4, torch.sin(u * th) / s.clamp_min(1e-8), torch.full_like(s, u))
        q = w0 * q0 + w1 * q1
        t = (1 - u) * keys[k]["global_trans"] + u * keys[k + 1]["global_trans"]
        out.append({"local_rotation": q.contiguous(), "global_trans": t.contiguous()})
    return out


def make_scene(N: int, J: int, width: int, height: int, n_views: int = 1, n_poses: Optional[int] = None,
               seed: int = 1234, isotropic: bool = False, soft_mask: bool = False,
               with_mlp_weight: bool = False, with_offsets: bool = False,
               pose_sequence: bool = False, scale_mu: float = 0.012, cam_radius: float = 4.0) -> Scene:
    g = _gen(seed)
    joints, parents = make_skeleton(J, g)
    bone = torch.randint(1, J, (N,), generator=g)
    s = torch.rand(N, 1, generator=g)
    a = joints[parents[bone]]
    b = joints[bone]
    xyz = a + s * (b - a) + 0.08 * torch.randn(N, 3, generator=g)
    scale = torch.exp(math.log(scale_mu) + 0.5 * torch.randn(N, 1 if isotropic else 3, generator=g))
    scaling_raw = torch.log(scale)
    rotation_raw = torch.randn(N, 4, generator=g)
    opac = 0.05 + 0.90 * torch.rand(N, 1, generator=g)
    opacity_raw = torch.log(opac / (1 - opac))
    features_dc = 0.5 * torch.randn(N, 1, 3, generator=g)
    features_rest = 0.1 * torch.randn(N, 15, 3, generator=g)
    if soft_mask:
        motion_mask = torch.sigmoid(2.0 * torch.randn(N, 1, generator=g))
    else:
        motion_mask = torch.ones(N, 1)
    scene_range = xyz.max() - xyz.min()
    node_radius_raw = torch.log(0.1 * scene_range + 1e-7) * torch.ones(J)
    node_radius_raw = node_radius_raw + 0.2 * torch.randn(J, generator=g)
    cams = make_cameras(n_views, width, height, g, radius=cam_radius)
    n_poses = n_views if n_poses is None else n_poses
    if pose_sequence:
        poses = slerp_pose_sequence(J, n_poses, g)
    else:
        poses = [make_pose(J, g) for _ in range(n_poses)]
    mlp_w = torch.sigmoid(torch.randn(N, J - 1, generator=g)) if with_mlp_weight else None
    offs = 0.01 * torch.randn(N, 3, generator=g) if with_offsets else None
    return Scene(joints.contiguous(), parents, node_radius_raw.contiguous(), xyz.contiguous(),
                 scaling_raw.contiguous(), rotation_raw.contiguous(), opacity_raw.contiguous(),
                 features_dc.contiguous(), features_rest.contiguous(), motion_mask.contiguous(), 3,
                 cams, poses, mlp_w, offs)


# The named configurations of BASELINE.json / SURVEY.md section 8(d).
CONFIGS = {
    "C0": dict(N=300_000, J=24, width=800, height=800, n_views=1, seed=1234),
    "C1": dict(N=100_000, J=24, width=800, height=800, n_views=1, seed=1235),
    "C2": dict(N=200_000, J=20, width=800, height=800, n_views=1, seed=1236),
    "C3": dict(N=300_000, J=24, width=1024, height=1024, n_views=4, seed=1237),
    "C4": dict(N=1_000_000, J=24, width=1920, height=1080, n_views=8, seed=1238),
    "C5": dict(N=2_000_000, J=60, width=800, height=800, n_views=1, n_poses=256, seed=1239,
               pose_sequence=True),
    # small parity cases
    "tiny": dict(N=4096, J=6, width=128, height=128, n_views=2, seed=7, cam_radius=1.2),
    "small": dict(N=20_000, J=24, width=256, height=192, n_views=2, seed=11, cam_radius=3.0),
    "ragged": dict(N=3001, J=7, width=97, height=61, n_views=1, seed=13, cam_radius=1.0, scale_mu=0.02),
}


def make_config(name: str, **overrides) -> Scene:
    kw = dict(CONFIGS[name])
    kw.update(overrides)
    return make_scene(**kw)


def seeded_cotangents(height: int, width: int, seed: int = 99):
    """Fixed upstream gradients for parity runs."""
    g = _gen(seed)
    hw = height * width
    return (torch.randn(3, height, width, generator=g) / (3 * hw),
            torch.randn(1, height, width, generator=g) / (10 * hw),
            torch.randn(1, height, width, generator=g) / (10 * hw))
