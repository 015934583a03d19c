"""CPU-only: the oracles against the golden vectors captured from the real reference.

* skinning: tests/golden/skin_*.npz, dq_blend.npz, sh_eval.npz come from importing the reference's Python
  (tests/golden/make_skin_golden.py, run in the build container);
* rasterizer: tests/golden/ref_cuda_*.npz come from the reference's CUDA kernels compiled unmodified and run
  on a B200 (tests/golden/make_ref_cuda_golden.py)."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import oracle_py as orc
from oracle import skin_oracle as so
from riggs_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLD, name + ".npz")).items()}


@pytest.mark.parametrize("name,dt,tol", [("skin_f64_j7", torch.float64, 1e-12), ("skin_f64_j24_plain", torch.float64, 1e-12),
                                         ("skin_f32_j24", torch.float32, 5e-6), ("skin_f32_j20_plain", torch.float32, 5e-6),
                                         ("skin_f64_j24_k3", torch.float64, 1e-12), ("skin_f32_j20_k4", torch.float32, 5e-6)])
def test_skin_oracle_vs_reference(name, dt, tol):
    d = _load(name)
    f = lambda k: d[k].to(dt) if k in d else None
    q = f("local_rotation").clone().requires_grad_(True)
    g = f("global_trans").clone().requires_grad_(True)
    nr = f("node_radius_raw").clone().requires_grad_(True)
    mask = f("motion_mask").clone().requires_grad_(True)
    s, o = f("mlp_weight"), f("template_offsets")
    if s is not None:
        s = s.clone().requires_grad_(True)
    if o is not None:
        o = o.clone().requires_grad_(True)
    K = d["nn_idx"].shape[1] if d["nn_idx"].shape[1] < d["joints"].shape[0] - 1 else -1  # K > 0: the nearest bones only
    out = so.deform_by_pose(f("xyz"), f("joints"), d["parents"], nr, q, g, mask, s, o, K=K)
    ((out["d_xyz"] * f("cot_xyz")).sum() + (out["d_rotation"] * f("cot_rot")).sum()).backward()
    for k in ("d_xyz", "d_rotation", "d_nodes", "nn_weight"):
        assert helpers.rel_err(out[k].detach().numpy(), d[k].numpy()) <= tol, k
    assert torch.equal(out["nn_idx"], d["nn_idx"])
    for k, v in (("g_local_rotation", q.grad), ("g_global_trans", g.grad), ("g_node_radius", nr.grad),
                 ("g_motion_mask", mask.grad), ("g_mlp_weight", None if s is None else s.grad),
                 ("g_template_offsets", None if o is None else o.grad)):
        if v is not None:
            assert helpers.rel_err(v.numpy(), d[k].numpy()) <= tol * 10, k


def test_dq_oracle_vs_reference():
    d = _load("dq_blend")
    R, T = so.dq_blend(d["q"].double(), d["t"].double(), d["w"].double())
    assert float((R - d["R"]).abs().max()) < 1e-12 and float((T - d["T"]).abs().max()) < 1e-12


def test_sh_oracle_vs_reference_eval_sh():
    """oracle preprocess SH->RGB against utils/sh_utils.eval_sh (golden), degrees 0..3."""
    d = _load("sh_eval")
    N = d["sh"].shape[0]
    cam = synth.look_at_camera(torch.tensor([0.0, -3.0, 0.0]), 64, 64, 0.7)
    campos = cam.camera_center
    means = campos[None] + d["dirs"] * 3.0  # direction from the camera = dirs; some land outside the frustum
    for deg in range(4):
        st = orc.forward(means.numpy(), np.full((N, 1), 0.5, np.float32), cam.world_view_transform.numpy(),
                         cam.full_proj_transform.numpy(), campos.numpy(), 64, 64, 0.36, 0.36, np.zeros(3, np.float32),
                         shs=d["sh"].numpy(), scales=np.full((N, 3), 0.01, np.float32),
                         rotations=np.tile(np.array([[1, 0, 0, 0]], np.float32), (N, 1)), degree=deg)
        vis = st["radii"] > 0
        assert vis.sum() >= 3
        expect = np.clip(d[f"rgb{deg}"].numpy() + 0.5, 0, None)
        assert np.allclose(st["rgb"][vis], expect[vis], atol=2e-6)
        assert np.array_equal(st["clamped"][vis].astype(bool), (d[f"rgb{deg}"].numpy() + 0.5 < 0)[vis])


REF_CASES = {"ragged": dict(config="ragged", bg=(0.2, 0.4, 0.6)), "mini": dict(config="tiny", bg=(0.0, 0.0, 0.0), N=700)}


@pytest.mark.parametrize("case", sorted(REF_CASES))
def test_raster_oracle_vs_reference_cuda_golden(case):
    path = os.path.join(GOLD, f"ref_cuda_{case}.npz")
    if not os.path.exists(path):
        pytest.skip("reference-CUDA fixture not captured yet (tests/golden/make_ref_cuda_golden.py on a B200)")
    ref = dict(np.load(path))
    c = REF_CASES[case]
    over = {k: v for k, v in c.items() if k not in ("config", "bg")}
    scene = synth.make_config(c["config"], **over)
    inp = helpers.raster_inputs(scene, view=0, bg=c["bg"])
    # (1) integer stages from the REFERENCE's per-Gaussian floats: bit-exact
    ov = dict(radii=ref["radii"], means2D=ref["means2D"], depths=ref["depths"], tiles_touched=ref["tiles_touched"],
              conic_opacity=ref["conic_opacity"], rgb=ref["rgb"])
    o = helpers.oracle_forward(inp, geom_override=ov)
    assert o["num_rendered"] == int(ref["num_rendered"])
    assert np.array_equal(o["keys"], ref["keys"]) and np.array_equal(o["point_list"], ref["point_list"])
    assert np.array_equal(o["ranges"], ref["ranges"])
    for k in ("color", "depth", "alpha"):
        assert helpers.rel_err(o[k], ref[k]) <= 1e-5, k
    assert (o["n_contrib"] != ref["n_contrib"]).mean() < 1e-3
    # (2) float stages on the CPU: rounding-level agreement, a handful of discrete flips at most
    o2 = helpers.oracle_forward(inp)
    assert (o2["radii"] != ref["radii"]).mean() < 2e-3
    both = (o2["radii"] > 0) & (ref["radii"] > 0)
    assert helpers.rel_err(o2["means2D"][both], ref["means2D"][both]) <= 1e-5
    assert helpers.rel_err(o2["conic_opacity"][both], ref["conic_opacity"][both]) <= 1e-4
    assert helpers.rel_err(o2["rgb"][both], ref["rgb"][both]) <= 1e-5
    assert np.array_equal(o2["clamped"][both], ref["clamped"][both])
    # (3) backward from the reference's forward state
    cots = synth.seeded_cotangents(inp["H"], inp["W"])
    o.update(cov3D=o2["cov3D"], clamped=ref["clamped"].astype(np.uint8), n_contrib=ref["n_contrib"], alpha=ref["alpha"])
    go = helpers.oracle_backward(o, inp, cots)
    same = (o2["radii"] > 0) == (ref["radii"] > 0)
    for k in ("dmean2D", "dcolor", "dopacity", "dmean3D", "dcov3D", "dsh", "dscale", "drot"):
        a, b = go[k], ref["g_" + k]
        P = a.shape[0]
        b = b.reshape(P, -1)
        if k == "dmean2D":
            b = b[:, :2]
        assert helpers.rel_err(a.reshape(P, -1)[same], b[same]) <= 1e-4, k


def test_oracle_integer_stage_properties():
    scene = synth.make_config("small")
    o = helpers.oracle_forward(helpers.raster_inputs(scene, view=1))
    keys = o["keys"]
    assert np.all(keys[1:] >= keys[:-1])
    assert np.array_equal(np.sort(o["keys_unsorted"], kind="stable"), keys)
    lens = o["ranges"][:, 1].astype(np.int64) - o["ranges"][:, 0]
    assert lens.sum() == o["num_rendered"] == int(o["tiles_touched"].sum())


def test_knn_oracle_small():
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3], [5, 5, 5]], np.float32)
    out = orc.dist2_knn3(pts)
    assert np.isclose(out[0], (1 + 4 + 9) / 3)
    d = ((pts[4] - pts[:4]) ** 2).sum(1)
    assert np.isclose(out[4], np.sort(d)[:3].mean())


@pytest.mark.parametrize("tag,tol", [("a", 1e-12), ("b", 2e-6), ("c", 1e-12), ("d", 1e-12)])
def test_loss_oracle_vs_reference_golden(tag, tol):
    """oracle/loss_oracle.py against outputs of the reference's utils/loss_utils.py (l1_loss, ssim, their autograd)."""
    from oracle import loss_oracle as lo
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_ssim.npz"))
    img = torch.from_numpy(g[f"{tag}_img"]).requires_grad_(True)
    gt = torch.from_numpy(g[f"{tag}_gt"])
    loss, l1, s = lo.image_loss(img, gt, float(g[f"{tag}_lambda"]))
    loss.backward()
    loss, l1, s = loss.detach(), l1.detach(), s.detach()
    assert abs(float(l1) - float(g[f"{tag}_l1"])) <= tol
    assert abs(float(s) - float(g[f"{tag}_ssim"])) <= tol
    assert abs(float(loss) - float(g[f"{tag}_loss"])) <= tol
    assert helpers.rel_err(img.grad.numpy(), g[f"{tag}_grad"]) <= max(tol, 1e-12) * 10


@pytest.mark.parametrize("name", ["node_warp_local", "node_warp_plain", "node_warp_k4"])
def test_node_warp_oracle_matches_reference_golden(name):
    """oracle.skin_oracle.node_warp against outputs and autograd gradients of the reference's ControlNodeWarp
    (tests/golden/make_node_warp_golden.py)."""
    g = {k: v for k, v in np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz")).items()}
    T = lambda k, rg=False: torch.from_numpy(g[k]).clone().requires_grad_(rg)
    K, hyper = int(g["K"]), int(g["hyper"])
    nodes, radius = T("nodes", True), T("node_radius_raw", True)
    nw = T("node_weight_raw", True) if int(g["with_node_weight"]) else None
    feature = T("feature", True) if "feature" in g else None
    attrs = {k[5:]: T(k, True) for k in g if k.startswith("node_d_") or k == "node_local_rotation"}
    mask = T("mask", True)
    out = so.node_warp(T("x"), feature, nodes, radius, nw, attrs, mask, K=K, hyper_dim=hyper,
                       local_frame=bool(g["local_frame"]), d_rot_as_res=bool(g["d_rot_as_res"]))
    assert np.array_equal(out["nn_idx"].numpy(), g["nn_idx"])
    np.testing.assert_allclose(out["nn_weight"].detach().numpy(), g["nn_weight"], rtol=1e-5, atol=1e-7)
    loss = 0
    for k in ("d_xyz", "d_rotation", "d_scaling", "d_opacity", "d_color"):
        if "out_" + k in g:
            np.testing.assert_allclose(out[k].detach().numpy(), g["out_" + k], rtol=1e-5, atol=1e-6)
            loss = loss + (out[k] * T("cot_" + k)).sum()
    loss.backward()
    close = lambda a, b: np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-5 * max(1.0, float(np.abs(b).max())))
    close(mask.grad.numpy(), g["g_mask"])
    close(radius.grad.numpy(), g["g_node_radius_raw"])
    if nw is not None:
        close(nw.grad.numpy(), g["g_node_weight_raw"])
    if feature is not None:
        close(feature.grad.numpy(), g["g_feature"])
        close(nodes.grad.numpy(), g["g_nodes"])
    for k, v in attrs.items():
        if v.grad is not None:
            close(v.grad.numpy(), g["g_node_" + k])


def test_dq_oracle_gradients_vs_reference_autograd():
    """skin_oracle.dq_blend differentiated by autograd against the gradients of the reference's own DQBlending
    (tests/golden/dq_blend_grad.npz, made by make_skin_golden.py --dq-grad)."""
    d = _load("dq_blend_grad")
    t = d["t"].double().clone().requires_grad_(True)
    w = d["w"].double().clone().requires_grad_(True)
    R, T = so.dq_blend(d["q"].double(), t, w)
    assert float((R - d["R"]).abs().max()) < 1e-12 and float((T - d["T"]).abs().max()) < 1e-12
    ((R * d["cR"].double()).sum() + (T * d["cT"].double()).sum()).backward()
    assert float((t.grad - d["g_t"]).abs().max()) < 1e-10
    assert float((w.grad - d["g_w"]).abs().max()) < 1e-10
