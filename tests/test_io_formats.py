"""CPU-only: the reference's on-disk formats (riggs_b200/io.py) -- property order, channel-major SH layout, round
trips, isotropic scaling, ASCII files."""
import os

import numpy as np
import torch

import helpers  # noqa: F401
from riggs_b200 import io as rio


def _model(P=37, fea=0, iso=False, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    return dict(xyz=r(P, 3), features_dc=r(P, 1, 3), features_rest=r(P, 15, 3), opacity=r(P, 1),
                scaling=r(P, 1 if iso else 3), rotation=r(P, 4), feature=r(P, fea) if fea else None)


def test_property_order_is_the_reference_s():
    names = rio.construct_list_of_attributes(fea_dim=2)
    # scene/gaussian_model.py:237-251
    want = (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(45)] +
            ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)] + ["fea_0", "fea_1"])
    assert names == want


def test_ply_round_trip_and_layout(tmp_path):
    m = _model(fea=2)
    path = os.path.join(tmp_path, "point_cloud", "iteration_1", "point_cloud.ply")
    rio.save_ply(path, **m)
    raw = open(path, "rb").read()
    header, body = raw.split(b"end_header\n", 1)
    lines = header.decode().split("\n")
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 37"]
    props = [l.split()[-1] for l in lines if l.startswith("property float")]
    assert props == rio.construct_list_of_attributes(fea_dim=2) and len(body) == 37 * 4 * len(props)
    table = np.frombuffer(body, "<f4").reshape(37, len(props))
    # channel-major SH: f_rest_k = features_rest[:, k % 15, k // 15] (transpose(1, 2).flatten(1), :258-259)
    assert np.array_equal(table[:, 9 + 16], m["features_rest"][:, 1, 1].numpy())
    assert np.array_equal(table[:, 3:6], np.zeros((37, 3), np.float32))
    back = rio.load_ply(path, fea_dim=2)
    for k, v in m.items():
        assert torch.equal(back[k], v), k


def test_ply_isotropic_and_ascii(tmp_path):
    m = _model(iso=True, seed=3)
    p = os.path.join(tmp_path, "iso.ply")
    rio.save_ply(p, **m)
    back = rio.load_ply(p, use_isotropic_gs=True)
    assert back["scaling"].shape == (37, 1) and torch.equal(back["scaling"], m["scaling"])
    # an ASCII file with the same header
    v = rio.read_ply_vertices(p)
    names = list(v)
    pa = os.path.join(tmp_path, "ascii.ply")
    with open(pa, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment test\nelement vertex 37\n" + "".join(f"property float {n}\n" for n in names) + "end_header\n")
        for i in range(37):
            f.write(" ".join(repr(float(v[n][i])) for n in names) + "\n")
    back2 = rio.load_ply(pa, use_isotropic_gs=True)
    for k in ("xyz", "features_rest", "rotation"):
        assert torch.allclose(back2[k], m[k], rtol=0, atol=1e-7), k


def test_skeleton_tree_round_trip(tmp_path):
    joints = torch.randn(24, 3)
    parents = torch.tensor([-1] + [max(0, i - 2) for i in range(1, 24)])
    p = os.path.join(tmp_path, "skeleton_tree.npz")
    rio.save_skeleton_tree(p, joints, parents, template_idx=5)
    z = np.load(p)
    assert sorted(z.files) == ["indices", "nodes", "parents", "template_idx"]  # train_rig.py:233
    t = rio.load_skeleton_tree(p)
    assert torch.equal(t["joints"], joints) and torch.equal(t["parents"], parents) and t["template_idx"] == 5
    np.savez(os.path.join(tmp_path, "mst.npz"), nodes=joints.numpy(), parents=np.zeros((24, 3)), children=np.zeros((24, 3)),
             parent_indices=parents.numpy())  # skeleton_utils/mst_utils.py:348
    assert torch.equal(rio.load_skeleton_tree(os.path.join(tmp_path, "mst.npz"))["parents"], parents)


def test_skeleton_state_dict_round_trip(tmp_path):
    from riggs_b200.skeleton_warp import SkeletonWarp
    joints = torch.randn(7, 3)
    parents = torch.tensor([-1, 0, 1, 1, 0, 4, 5])
    a = SkeletonWarp(joints, parents, device="cpu")
    b = SkeletonWarp(torch.zeros(7, 3), parents, device="cpu")
    with torch.no_grad():
        a._node_radius.add_(0.3)
    p = os.path.join(tmp_path, "skeleton.pth")
    torch.save(rio.skeleton_state_dict(a), p)  # scene/skeleton_model.py:46
    state = torch.load(p)
    assert "nodes" in state and "_node_radius" in state and any(k.startswith("pose_net") for k in state)
    assert rio.load_skeleton_state_dict(b, {**state, "gs__xyz": torch.zeros(3)}) == ["gs__xyz"]
    assert torch.equal(b.nodes, a.nodes) and torch.equal(b._node_radius, a._node_radius)
    for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        assert ka == kb and torch.equal(va, vb), ka


def test_weight_mlp_cache_is_keyed_on_the_cloud():
    """SkeletonWarp._weight_mlp_output: one evaluation per canonical cloud under no_grad, none cached under autograd."""
    from riggs_b200.skeleton_warp import SkeletonWarp
    m = SkeletonWarp(torch.randn(5, 3), torch.tensor([-1, 0, 1, 1, 0]), device="cpu")
    calls = []
    orig = m.skinning_weight_mlp.forward
    m.skinning_weight_mlp.forward = lambda x: (calls.append(1), orig(x))[1]
    x = torch.randn(40, 3)
    m.cache_skinning_weights = True
    with torch.no_grad():
        a = m._weight_mlp_output(x)
        b = m._weight_mlp_output(x)
        assert len(calls) == 1 and a is b
        x.add_(1.0)  # in-place change of the cloud: the cache must not be used
        c = m._weight_mlp_output(x)
        assert len(calls) == 2 and not torch.equal(a, c)
        m.clear_weight_cache()
        m._weight_mlp_output(x)
        assert len(calls) == 3
    m._weight_mlp_output(x)  # autograd on: evaluated, not cached
    m._weight_mlp_output(x)
    assert len(calls) == 5
    m.cache_skinning_weights = False
    with torch.no_grad():
        m._weight_mlp_output(x)
        m._weight_mlp_output(x)
    assert len(calls) == 7
