"""Generate tests/golden/skin_*.npz by running the REAL reference skinning code on the CPU.

Run in the build container only (needs /root/reference):

    python tests/golden/make_skin_golden.py

It imports ``skeleton_utils.skeleton_warp.SkeletonWarp`` and ``utils.dual_quaternion`` from the
read-only reference tree with the two import shims of SURVEY.md section 8(c) (stub ``pytorch3d``,
identity ``.cuda()``), feeds them seeded synthetic inputs from ``riggs_b200.synth`` and stores the
inputs, the outputs and the autograd gradients.  The MLP front ends (WeightMLP / detail_net) are
replaced by fixed tensors so that the golden data pins the LBS core, which is what the CUDA
kernels implement; the reference's ``deform_by_pose`` body itself runs unmodified.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
p3d, ops = types.ModuleType("pytorch3d"), types.ModuleType("pytorch3d.ops")
ops.knn_points = ops.ball_query = None
p3d.ops = ops
sys.modules["pytorch3d"], sys.modules["pytorch3d.ops"] = p3d, ops
torch.Tensor.cuda = lambda self, *a, **k: self
torch.nn.Module.cuda = lambda self, *a, **k: self

from skeleton_utils.skeleton_warp import SkeletonWarp  # noqa: E402
from utils import dual_quaternion as refdq  # noqa: E402
from utils.sh_utils import eval_sh  # noqa: E402

from riggs_b200 import synth  # noqa: E402


class _Fixed(torch.nn.Module):
    def __init__(self, t):
        super().__init__()
        self.t = t

    def forward(self, *a, **k):
        return self.t


def one_case(name, N, J, seed, dtype, mlp, offs, soft_mask, K=-1):
    sc = synth.make_scene(N=N, J=J, width=64, height=64, n_views=1, seed=seed, soft_mask=soft_mask,
                          with_mlp_weight=mlp, with_offsets=offs)
    m = SkeletonWarp(is_blender=True, joints=sc.joints, parent_indices=sc.parents, K=K,
                     use_skinning_weight_mlp=True, use_template_offsets=True)
    m = m.to(dtype)
    m._node_radius.data = sc.node_radius_raw.to(dtype).clone()
    x = sc.xyz.to(dtype)
    q = sc.poses[0]["local_rotation"].to(dtype).clone().requires_grad_(True)
    g = sc.poses[0]["global_trans"].to(dtype).clone().requires_grad_(True)
    mask = sc.motion_mask.to(dtype).clone().requires_grad_(True)
    m.use_skinning_weight_mlp = mlp
    m.use_template_offsets = offs
    s = o = None
    if mlp:
        s = sc.mlp_weight.to(dtype).clone().requires_grad_(True)
        m.skinning_weight_mlp = _Fixed(s)
    if offs:
        o = sc.template_offsets.to(dtype).clone().requires_grad_(True)
        m.detail_net = _Fixed(o)
    out = m.deform_by_pose(x, {"local_rotation": q, "global_trans": g}, mask)
    gen = torch.Generator().manual_seed(seed + 100)
    a = torch.randn(N, 3, generator=gen).to(dtype)
    h = torch.randn(N, 4, generator=gen).to(dtype)
    loss = (out["d_xyz"] * a).sum() + (out["d_rotation"] * h).sum()
    loss.backward()
    d = dict(joints=sc.joints, parents=sc.parents, node_radius_raw=sc.node_radius_raw, xyz=sc.xyz,
             local_rotation=sc.poses[0]["local_rotation"], global_trans=sc.poses[0]["global_trans"],
             motion_mask=sc.motion_mask, cot_xyz=a, cot_rot=h,
             d_xyz=out["d_xyz"].detach(), d_rotation=out["d_rotation"].detach(),
             d_nodes=out["d_nodes"].detach(), nn_weight=out["nn_weight"].detach(),
             nn_idx=out["nn_idx"], g_local_rotation=q.grad, g_global_trans=g.grad,
             g_node_radius=m._node_radius.grad, g_motion_mask=mask.grad)
    if mlp:
        d["mlp_weight"] = sc.mlp_weight
        d["g_mlp_weight"] = s.grad
    if offs:
        d["template_offsets"] = sc.template_offsets
        d["g_template_offsets"] = o.grad
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        **{k: (v.detach().numpy() if torch.is_tensor(v) else v) for k, v in d.items()})
    print("wrote", name, {k: tuple(v.shape) for k, v in d.items() if torch.is_tensor(v)})


def dq_case():
    gen = torch.Generator().manual_seed(5)
    K, N = 7, 40
    q = torch.randn(K, 4, generator=gen)  # un-normalised on purpose: QT2DQ normalises (dual_quaternion.py:138)
    t = torch.randn(K, 3, generator=gen)
    w = torch.softmax(torch.randn(N, K, generator=gen), -1)
    # Call shape q (K,4), t (K,3), weights (N,K): with it F.normalize's default dim=1 is the quaternion
    # axis.  (transformation_blending (:189-197) passes q[None], for which dim=1 is the bone axis -- a
    # quirk of that dead code we do not reproduce.)
    R, T = refdq.DQBlending(q.double(), t.double(), w.double(), rot_as_q=False)
    qq, _ = refdq.DQBlending(q.double(), t.double(), w.double(), rot_as_q=True)
    np.savez_compressed(os.path.join(HERE, "dq_blend.npz"), q=q.numpy(), t=t.numpy(), w=w.numpy(),
                        R=R.numpy(), T=T.numpy(), q_out=qq.numpy())
    print("wrote dq_blend")


def dq_grad_case():
    """Gradients of the reference's DQBlending by its own autograd: dL/dt (bone translations) and dL/dweights for seeded
    cotangents on the blended rotation matrix and translation (pins the dual-quaternion backward of the kernels)."""
    gen = torch.Generator().manual_seed(15)
    K, N = 9, 33
    q = torch.randn(K, 4, generator=gen).double()
    t = torch.randn(K, 3, generator=gen).double().requires_grad_(True)
    w = torch.softmax(torch.randn(N, K, generator=gen), -1).double().requires_grad_(True)
    cR = torch.randn(N, 3, 3, generator=gen).double()
    cT = torch.randn(N, 3, generator=gen).double()
    R, T = refdq.DQBlending(q, t, w, rot_as_q=False)
    ((R * cR).sum() + (T * cT).sum()).backward()
    np.savez_compressed(os.path.join(HERE, "dq_blend_grad.npz"), q=q.numpy(), t=t.detach().numpy(), w=w.detach().numpy(),
                        cR=cR.numpy(), cT=cT.numpy(), R=R.detach().numpy(), T=T.detach().numpy(), g_t=t.grad.numpy(),
                        g_w=w.grad.numpy())
    print("wrote dq_blend_grad")


def sh_case():
    gen = torch.Generator().manual_seed(6)
    N = 64
    sh = torch.randn(N, 16, 3, generator=gen) * 0.3
    dirs = torch.nn.functional.normalize(torch.randn(N, 3, generator=gen), dim=-1)
    out = {}
    for deg in range(4):
        out[f"rgb{deg}"] = eval_sh(deg, sh.transpose(1, 2).double(), dirs.double()).numpy()
    np.savez_compressed(os.path.join(HERE, "sh_eval.npz"), sh=sh.numpy(), dirs=dirs.numpy(), **out)
    print("wrote sh_eval")


if __name__ == "__main__":
    torch.manual_seed(0)
    if "--dq-grad" in sys.argv:
        dq_grad_case()
        sys.exit(0)
    if "--topk" in sys.argv:  # only the K > 0 cases (added later; the files below are left untouched)
        # K > 0 with the weight MLP indexes the MLP output with bone numbers 1..J-1 (skeleton_warp.py:59) and fails
        # in the reference whenever the last bone is among the nearest: these cases run with the MLP off
        one_case("skin_f64_j24_k3", 300, 24, 25, torch.float64, False, True, True, K=3)
        one_case("skin_f32_j20_k4", 257, 20, 26, torch.float32, False, False, False, K=4)
        sys.exit(0)
    one_case("skin_f64_j7", 50, 7, 21, torch.float64, True, True, True)
    one_case("skin_f64_j24_plain", 200, 24, 22, torch.float64, False, False, False)
    one_case("skin_f32_j24", 300, 24, 23, torch.float32, True, True, True)
    one_case("skin_f32_j20_plain", 257, 20, 24, torch.float32, False, False, False)
    dq_case()
    sh_case()
