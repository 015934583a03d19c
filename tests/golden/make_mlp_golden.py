"""Generate tests/golden/mlp_front_end.npz by running the REAL reference MLP front ends on the CPU.

Run in the build container only (needs /root/reference):

    python tests/golden/make_mlp_golden.py

Imports ``skeleton_utils.network_utils.{PoseMLP, WeightMLP, DeformMLP}`` and ``SkeletonWarp`` from the read-only
reference tree with the import shims of SURVEY.md section 8(c), draws seeded weights, and stores
  * every parameter under the key it has in the reference's ``skeleton.pth`` (SkeletonWarp.state_dict():
    ``pose_net.net.{i}.weight``, ``skinning_weight_mlp.linear.{i}.weight``, ``detail_net.linear.{i}.weight`` ...),
  * seeded inputs and the reference modules' outputs,
  * the full ``SkeletonWarp.forward`` result (PoseMLP -> FK -> WeightMLP -> detail_net -> LBS) for one time stamp,
  * the complete list of state-dict keys and shapes of the reference's SkeletonWarp.
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

from riggs_b200 import synth  # noqa: E402


def main():
    torch.manual_seed(4242)
    J, N = 7, 257
    sc = synth.make_scene(N=N, J=J, width=64, height=64, n_views=1, seed=77)
    m = SkeletonWarp(is_blender=True, joints=sc.joints, parent_indices=sc.parents, K=-1,
                     use_skinning_weight_mlp=True, use_template_offsets=True)
    m._node_radius.data = sc.node_radius_raw.clone()
    with torch.no_grad():  # the reference initialises the offset head at 1e-5: give it something to show
        m.detail_net.gaussian_warp.weight.normal_(0, 0.05)
        m.detail_net.gaussian_warp.bias.normal_(0, 0.01)
        for p in m.pose_net.parameters():
            p.add_(0.02 * torch.randn_like(p))
    out = {}
    keys, shapes = [], []
    for k, v in m.state_dict().items():
        keys.append(k)
        shapes.append(",".join(str(s) for s in v.shape))
        if k.split(".")[0] in ("pose_net", "skinning_weight_mlp", "detail_net") or k in ("nodes", "_node_radius"):
            out["state/" + k] = v.detach().numpy()
    out["state_keys"] = np.array(keys)
    out["state_shapes"] = np.array(shapes)
    g = torch.Generator().manual_seed(9)
    x = sc.xyz.clone()
    t = torch.rand(1, generator=g)
    pose = torch.randn(N, 4 * J, generator=g) * 0.3
    with torch.no_grad():
        pm = m.pose_net(t)
        out.update(x=x.numpy(), t=t.numpy(), pose=pose.numpy(), pose_rotation=pm["rotation"].numpy(),
                   pose_translation=pm["translation"].numpy(), weight_mlp=m.skinning_weight_mlp(x).numpy(),
                   detail_net=m.detail_net(x, pose).numpy())
        full = m.forward(x, t.reshape(1, 1).expand(J, 1), sc.motion_mask)
        out.update(fwd_d_xyz=full["d_xyz"].numpy(), fwd_d_rotation=full["d_rotation"].numpy(),
                   fwd_nn_weight=full["nn_weight"].numpy(), fwd_local_rotation=full["local_rotation"].numpy(),
                   fwd_global_trans=full["global_trans"].numpy(), fwd_d_nodes=full["d_nodes"].numpy())
    out.update(joints=sc.joints.numpy(), parents=sc.parents.numpy(), motion_mask=sc.motion_mask.numpy())
    path = os.path.join(HERE, "mlp_front_end.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path) >> 10, "KiB;", len(keys), "state keys")
    for k, s in zip(keys, shapes):
        print("  ", k, s)


if __name__ == "__main__":
    main()
