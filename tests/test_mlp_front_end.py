"""The torch MLP front ends of the skinning path (PoseMLP, WeightMLP, DeformMLP = detail_net) against the REFERENCE's own
modules: tests/golden/mlp_front_end.npz holds seeded weights under the reference's state-dict keys, seeded inputs and
the outputs of skeleton_utils/network_utils.py run on the CPU (tests/golden/make_mlp_golden.py).  Loading is by key, so
these tests also pin that a reference ``skeleton.pth`` fits the modules here."""
import os

import numpy as np
import pytest
import torch

import helpers  # noqa: F401
from riggs_b200 import io as rio
from riggs_b200 import skeleton_warp as sw

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "mlp_front_end.npz"))
STATE = {k[len("state/"):]: torch.from_numpy(G[k]) for k in G.files if k.startswith("state/")}
J = int(G["joints"].shape[0])


def _sub(prefix):
    return {k[len(prefix) + 1:]: v for k, v in STATE.items() if k.startswith(prefix + ".")}


def test_pose_mlp_matches_reference():
    m = sw.PoseMLP(1, 4 * J)
    m.load_state_dict(_sub("pose_net"), strict=True)
    with torch.no_grad():
        out = m(torch.from_numpy(G["t"]))
    assert np.allclose(out["rotation"].numpy(), G["pose_rotation"], rtol=0, atol=1e-6)
    assert np.allclose(out["translation"].numpy(), G["pose_translation"], rtol=0, atol=1e-6)


def test_weight_mlp_matches_reference():
    m = sw.WeightMLP(3, J - 1)
    m.load_state_dict(_sub("skinning_weight_mlp"), strict=True)
    with torch.no_grad():
        out = m(torch.from_numpy(G["x"]))
    assert out.shape == (G["x"].shape[0], J - 1)
    assert np.allclose(out.numpy(), G["weight_mlp"], rtol=0, atol=1e-6)


def test_detail_net_matches_reference():
    m = sw.DeformMLP(4 * J)
    m.load_state_dict(_sub("detail_net"), strict=True)
    with torch.no_grad():
        out = m(torch.from_numpy(G["x"]), torch.from_numpy(G["pose"]))
    assert np.allclose(out.numpy(), G["detail_net"], rtol=0, atol=1e-6)


def test_reference_skeleton_state_dict_loads_by_key():
    """Every parameter of the three MLPs, ``nodes`` and ``_node_radius`` of a reference skeleton.pth find their
    counterpart; what is reported as unknown is only what this path does not use (the stage-1 network the reference's
    SkeletonWarp inherits from ControlNodeWarp, its 512 control nodes, node weights)."""
    warp = sw.SkeletonWarp(torch.from_numpy(G["joints"]), torch.from_numpy(G["parents"]), device="cpu")
    full = {str(k): torch.zeros([int(x) for x in str(s).split(",")] if str(s) else []) for k, s in
            zip(G["state_keys"], G["state_shapes"])}
    full.update(STATE)
    unknown = rio.load_skeleton_state_dict(warp, full)
    used = ("pose_net.", "skinning_weight_mlp.", "detail_net.")
    assert not [k for k in unknown if k.startswith(used) or k in ("nodes", "_node_radius")], unknown
    assert all(k.startswith("network.") or k in ("control_nodes", "inited", "_node_weight", "gs", "init_pcl") or "node" in k for k in unknown), unknown
    for k, v in STATE.items():
        assert torch.equal(warp.state_dict()[k], v), k
    with pytest.raises(KeyError):
        rio.load_skeleton_state_dict(warp, {"pose_net.net.0.weight": torch.zeros(3, 3)}, strict=True)


@pytest.mark.gpu
def test_skeleton_warp_forward_matches_reference_forward():
    """SkeletonWarp.forward end to end (PoseMLP -> FK kernel -> WeightMLP -> detail_net -> LBS kernel) against the
    reference's own SkeletonWarp.forward with the same weights."""
    warp = sw.SkeletonWarp(torch.from_numpy(G["joints"]), torch.from_numpy(G["parents"]), device="cuda")
    assert rio.load_skeleton_state_dict(warp, STATE) == []
    x = torch.from_numpy(G["x"]).cuda()
    t = torch.from_numpy(G["t"]).cuda().reshape(1, 1).expand(J, 1)
    with torch.no_grad():
        out = warp(x, t, torch.from_numpy(G["motion_mask"]).cuda())
    for k in ("d_xyz", "d_rotation", "nn_weight", "local_rotation", "global_trans", "d_nodes"):
        assert helpers.rel_err(out[k].cpu().numpy(), G["fwd_" + k], floor=1e-3) <= 1e-5, k
