"""Generate tests/golden/node_warp_*.npz by running the reference's stage-1 control-node warp on the CPU.

Run in the build container only (needs /root/reference):

    python tests/golden/make_node_warp_golden.py

``utils.time_utils.ControlNodeWarp.forward`` / ``cal_nn_weight`` (utils/time_utils.py:934-967, 1133-1236) run
unmodified on a module object created without its constructor (the constructor builds the deformation MLP and
samples nodes from a point cloud on the GPU); the per-node motion the MLP would predict (``node_deform``) is
replaced by fixed seeded tensors, and ``pytorch3d.ops.knn_points`` -- a dependency that is not part of the
reference tree -- by a brute-force search with the same contract (squared distances, ascending).
Stored: inputs, outputs and the autograd gradients of a seeded linear functional of the outputs.
"""
import os
import sys
import types
from collections import namedtuple

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
torch.Tensor.cuda = lambda self, *a, **k: self
torch.nn.Module.cuda = lambda self, *a, **k: self

_KNN = namedtuple("KNN", "dists idx knn")


def knn_points(p1, p2, lengths1=None, lengths2=None, K=1, **kw):
    d = ((p1[:, :, None, :] - p2[:, None, :, :]) ** 2).sum(-1)  # differentiable, like pytorch3d's
    idx = torch.topk(d.detach().double(), K, dim=-1, largest=False, sorted=True).indices
    return _KNN(torch.gather(d, 2, idx), idx, None)


p3d, ops = types.ModuleType("pytorch3d"), types.ModuleType("pytorch3d.ops")
ops.knn_points, ops.ball_query = knn_points, None
p3d.ops = ops
sys.modules["pytorch3d"], sys.modules["pytorch3d.ops"] = p3d, ops

from utils.time_utils import ControlNodeWarp  # noqa: E402


def one_case(name, N, M, K, hyper, seed, local_frame, d_rot_as_res, with_node_weight, soft_mask, pred):
    g = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=g)
    m = ControlNodeWarp.__new__(ControlNodeWarp)
    torch.nn.Module.__init__(m)
    m.K, m.hyper_dim, m.skinning, m.local_frame, m.d_rot_as_res = K, hyper, False, local_frame, d_rot_as_res
    m.with_node_weight, m.pred_opacity, m.pred_color = with_node_weight, pred, pred
    m.cached_nn_weight, m.nn_weight = False, None
    m.lambda_arap_landmarks, m.lambda_arap_steps = [0], [0]
    x = rn(N, 3) * 0.5
    nodes = torch.cat([x[torch.randperm(N, generator=g)[:M]] + 0.02 * rn(M, 3),
                       1e-2 * torch.ones(M, hyper) + 0.01 * rn(M, hyper)], -1)
    m.nodes = torch.nn.Parameter(nodes.clone())
    m._node_radius = torch.nn.Parameter(torch.log(torch.tensor(0.25)) + 0.2 * rn(M))
    m._node_weight = torch.nn.Parameter(0.5 * rn(M, 1))
    feature = (1e-2 * torch.ones(N, hyper) + 0.01 * rn(N, hyper)).requires_grad_(True) if hyper > 0 else None
    attrs = {"d_xyz": (0.1 * rn(M, 3)).requires_grad_(True), "d_rotation": (0.1 * rn(M, 4)).requires_grad_(True),
             "d_scaling": (0.05 * rn(M, 3)).requires_grad_(True), "local_rotation": (0.2 * rn(M, 4)).requires_grad_(True)}
    if pred:
        attrs["d_opacity"] = (0.1 * rn(M, 1)).requires_grad_(True)
        attrs["d_color"] = (0.1 * rn(M, 3)).requires_grad_(True)
    m.node_deform = lambda t, **kw: dict(attrs)
    mask = (torch.sigmoid(2 * rn(N, 1)) if soft_mask else torch.ones(N, 1)).requires_grad_(True)
    m.eval()
    out = m.forward(x, torch.zeros(M, 1), feature, mask, is_training=False)
    nn_weight, nn_dist, nn_idx = m.cal_nn_weight(x=x, feature=feature)
    cot = {"d_xyz": rn(N, 3), "d_rotation": rn(N, 4), "d_scaling": rn(N, 3)}
    if pred:
        cot["d_opacity"], cot["d_color"] = rn(N, 1), rn(N, 3)
    loss = sum((out[k] * c).sum() for k, c in cot.items())
    loss.backward()
    d = dict(x=x, nodes=nodes, node_radius_raw=m._node_radius.detach(), node_weight_raw=m._node_weight.detach(),
             mask=mask.detach(), K=np.int64(K), hyper=np.int64(hyper), local_frame=np.int64(local_frame),
             d_rot_as_res=np.int64(d_rot_as_res), with_node_weight=np.int64(with_node_weight),
             nn_idx=nn_idx, nn_weight=nn_weight.detach(), nn_dist=nn_dist.detach(), d_nodes=out["d_nodes"].detach(),
             g_mask=mask.grad, g_node_radius_raw=m._node_radius.grad,
             g_nodes=m.nodes.grad if m.nodes.grad is not None else torch.zeros_like(nodes))
    if with_node_weight:
        d["g_node_weight_raw"] = m._node_weight.grad
    if feature is not None:
        d["feature"], d["g_feature"] = feature.detach(), feature.grad
    for k, v in attrs.items():
        d["node_" + k] = v.detach()
        d["g_node_" + k] = v.grad if v.grad is not None else torch.zeros_like(v)
    for k, c in cot.items():
        d["out_" + k] = out[k].detach()
        d["cot_" + k] = c
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        **{k: (v.numpy() if torch.is_tensor(v) else v) for k, v in d.items()})
    print(name, "ok", {k: tuple(v.shape) for k, v in out.items() if torch.is_tensor(v)})


if __name__ == "__main__":
    one_case("node_warp_local", N=1500, M=64, K=3, hyper=2, seed=11, local_frame=True, d_rot_as_res=True,
             with_node_weight=True, soft_mask=True, pred=False)
    one_case("node_warp_plain", N=1000, M=40, K=3, hyper=0, seed=12, local_frame=False, d_rot_as_res=False,
             with_node_weight=False, soft_mask=False, pred=True)
    one_case("node_warp_k4", N=800, M=100, K=4, hyper=2, seed=13, local_frame=True, d_rot_as_res=True,
             with_node_weight=True, soft_mask=False, pred=True)
