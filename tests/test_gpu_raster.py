"""GPU parity of the rasterizer: libriggs_b200 (through the C ABI) against
  (1) the reference's own CUDA kernels compiled unmodified (oracle/_ref) -- bit-exact gates + tolerances,
  (2) the plain-C CPU oracle.
Tolerances are the north star's: tile keys / sort order / ranges bit-exact, images 1e-5 relative,
gradients 1e-4 relative (relative to the largest magnitude of the reference array)."""
import numpy as np
import pytest
import torch

import helpers
from oracle import ref_cuda
from riggs_b200 import synth

pytestmark = pytest.mark.gpu

IMG_RTOL = 1e-5
GRAD_RTOL = 1e-4

CASES = [
    ("tiny", dict(), (0.0, 0.0, 0.0)),
    ("ragged", dict(), (0.2, 0.4, 0.6)),
    ("small", dict(), (1.0, 1.0, 1.0)),
    ("small", dict(isotropic=True, soft_mask=True, with_mlp_weight=True, with_offsets=True), (0.0, 0.0, 0.0)),
]


def _run_ref(cin, cots=None):
    ref = ref_cuda.RefRasterizer()
    f = ref.forward(cin["bg"], cin["means3D"], cin["opacities"], cin["view"], cin["proj"], cin["campos"], cin["W"],
                    cin["H"], cin["tanfovx"], cin["tanfovy"], shs=cin.get("shs"),
                    colors_precomp=cin.get("colors_precomp"), scales=cin.get("scales"), rotations=cin.get("rotations"),
                    cov3D_precomp=cin.get("cov3D_precomp"), degree=cin.get("degree", 3))
    st = ref.state_arrays()
    g = None
    if cots is not None:
        g = ref.backward(cin["bg"], cin["means3D"], f["radii"], f["alpha"], cin["view"], cin["proj"], cin["campos"],
                         cin["tanfovx"], cin["tanfovy"], cots[0], cots[1], cots[2], shs=cin.get("shs"),
                         colors_precomp=cin.get("colors_precomp"), scales=cin.get("scales"),
                         rotations=cin.get("rotations"), cov3D_precomp=cin.get("cov3D_precomp"),
                         degree=cin.get("degree", 3))
    return f, st, g


# Forward-only cases that stress the binning: 8160 tiles (1080p; groups of eight chunks at the shared-memory limit),
# rectangles far wider than a warp (large splats: the general and the warp-cooperative paths), more than 4096 chunks'
# worth of Gaussians (chunks longer than 128), and almost everything culled.
BINNING_CASES = [
    ("C4", dict(N=60_000, n_views=1), (0.0, 0.0, 0.0)),
    ("C4", dict(N=20_000, n_views=1, scale_mu=0.15), (0.0, 0.0, 0.0)),  # 1080p: rectangles far wider than a warp, cut by the bands
    ("small", dict(N=30_000, scale_mu=0.12, n_views=1), (0.0, 0.0, 0.0)),
    ("small", dict(N=600_000, n_views=1, scale_mu=0.004), (0.0, 0.0, 0.0)),
    ("tiny", dict(cam_radius=0.05, n_views=1), (0.0, 0.0, 0.0)),
]

needs_ref = pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref/libref_riggs.so not built")


@needs_ref
@pytest.mark.parametrize("name,over,bg", CASES + BINNING_CASES)
def test_forward_vs_reference_cuda(name, over, bg):
    scene = synth.make_config(name, **over)
    for view in range(len(scene.cameras)):
        cin = helpers.to_cuda(helpers.raster_inputs(scene, view=view, bg=bg))
        ours = helpers.ours_forward(cin)
        f, st, _ = _run_ref(cin)
        # ---- bit-exact gates
        assert ours["num_rendered"] == f["num_rendered"]
        assert np.array_equal(ours["radii"], f["radii"].cpu().numpy())
        assert np.array_equal(ours["tiles_touched"], st["tiles_touched"])
        vis = ours["radii"] > 0
        assert np.array_equal(ours["depths"][vis].view(np.uint32), st["depths"][vis].view(np.uint32))
        assert np.array_equal(ours["means2D"][vis].view(np.uint32), st["means2D"][vis].view(np.uint32))
        # our kernels never build the 64-bit keys; the sorted keys are rebuilt from (ranges, point_list, depths)
        assert np.array_equal(ours["keys"], st["keys"])
        assert np.array_equal(ours["point_list"], st["point_list"])
        assert np.array_equal(ours["ranges"], st["ranges"])
        assert np.array_equal(ours["n_contrib"], st["n_contrib"])
        # ---- floats
        assert helpers.rel_err(ours["conic_opacity"][vis], st["conic_opacity"][vis]) <= 1e-6
        assert helpers.rel_err(ours["rgb"][vis], st["rgb"][vis]) <= 1e-6
        for k in ("color", "depth", "alpha"):
            assert helpers.rel_err(ours[k].cpu().numpy(), f[k].cpu().numpy()) <= IMG_RTOL, k


@needs_ref
@pytest.mark.parametrize("name,over,bg", CASES)
def test_backward_vs_reference_cuda(name, over, bg):
    scene = synth.make_config(name, **over)
    cin = helpers.to_cuda(helpers.raster_inputs(scene, view=0, bg=bg))
    cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
    ours = helpers.ours_forward(cin)
    g = helpers.ours_backward(cin, ours, cots)
    _, _, gr = _run_ref(cin, cots)
    for k in ("dmean2D", "dcolor", "dopacity", "dmean3D", "dcov3D", "dsh", "dscale", "drot"):
        assert helpers.rel_err(g[k].cpu().numpy().reshape(-1), gr[k].cpu().numpy().reshape(-1)) <= GRAD_RTOL, k


@pytest.mark.parametrize("name,over,bg", CASES[:3])
def test_forward_backward_vs_cpu_oracle(name, over, bg):
    scene = synth.make_config(name, **over)
    inp = helpers.raster_inputs(scene, view=0, bg=bg)
    cin = helpers.to_cuda(inp)
    ours = helpers.ours_forward(cin)
    # Pure-integer stages must be bit-exact when the oracle starts from the same per-Gaussian floats.
    ov = dict(radii=ours["radii"], means2D=ours["means2D"], depths=ours["depths"], tiles_touched=ours["tiles_touched"],
              conic_opacity=ours["conic_opacity"], rgb=ours["rgb"])
    o = helpers.oracle_forward(inp, geom_override=ov)
    assert o["num_rendered"] == ours["num_rendered"]
    for k in ("keys", "point_list", "ranges"):
        assert np.array_equal(o[k], ours[k]), k
    for k in ("color", "depth", "alpha"):
        assert helpers.rel_err(ours[k].cpu().numpy(), o[k]) <= IMG_RTOL, k
    assert (o["n_contrib"] != ours["n_contrib"]).mean() < 1e-3  # expf differs in the last ulp between CPU and GPU
    # Float stages from scratch on the CPU: discrete outputs may flip for a handful of Gaussians (no FMA on CPU).
    o2 = helpers.oracle_forward(inp)
    assert (o2["radii"] != ours["radii"]).mean() < 2e-3
    both = (o2["radii"] > 0) & (ours["radii"] > 0)
    assert helpers.rel_err(ours["means2D"][both], o2["means2D"][both]) <= 1e-5
    assert helpers.rel_err(ours["conic_opacity"][both], o2["conic_opacity"][both]) <= 1e-4
    assert helpers.rel_err(ours["rgb"][both], o2["rgb"][both]) <= 1e-5
    # Backward against the CPU oracle run on OUR forward state.
    cots = synth.seeded_cotangents(inp["H"], inp["W"])
    g = helpers.ours_backward(cin, ours, cots)
    o.update(cov3D=o2["cov3D"], clamped=np.stack([(ours["clamped_bits"] >> c) & 1 for c in range(3)], 1).astype(np.uint8),
             n_contrib=ours["n_contrib"], alpha=ours["alpha"].cpu().numpy())
    go = helpers.oracle_backward(o, inp, cots)
    same = (o2["radii"] > 0) == (ours["radii"] > 0)  # rows whose cull decision agrees (cov3D of the others is unset)
    for k in ("dmean2D", "dcolor", "dopacity", "dmean3D", "dcov3D", "dsh", "dscale", "drot"):
        a = g[k].cpu().numpy()
        b = go[k]
        if k == "dmean2D":
            a = a[:, :2]
        P = a.shape[0]
        assert helpers.rel_err(a.reshape(P, -1)[same], b.reshape(P, -1)[same]) <= GRAD_RTOL, k


@needs_ref
def test_precomputed_paths_vs_reference_cuda():
    """colors_precomp and cov3D_precomp select the other code paths (null-pointer convention)."""
    scene = synth.make_config("tiny")
    inp = helpers.raster_inputs(scene, view=1, bg=(0.1, 0.2, 0.3))
    o = helpers.oracle_forward(inp)
    g = torch.Generator().manual_seed(3)
    inp2 = dict(inp)
    inp2.pop("shs"); inp2.pop("scales"); inp2.pop("rotations")
    inp2["colors_precomp"] = torch.rand(inp["means3D"].shape[0], 3, generator=g)
    inp2["cov3D_precomp"] = torch.from_numpy(o["cov3D"].copy())
    cin = helpers.to_cuda(inp2)
    cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
    ours = helpers.ours_forward(cin)
    gg = helpers.ours_backward(cin, ours, cots)
    f, st, gr = _run_ref(cin, cots)
    assert np.array_equal(ours["keys"], st["keys"]) and np.array_equal(ours["point_list"], st["point_list"])
    assert np.array_equal(ours["ranges"], st["ranges"])
    for k in ("color", "depth", "alpha"):
        assert helpers.rel_err(ours[k].cpu().numpy(), f[k].cpu().numpy()) <= IMG_RTOL, k
    for k in ("dmean2D", "dcolor", "dopacity", "dmean3D", "dcov3D"):
        assert helpers.rel_err(gg[k].cpu().numpy().reshape(-1), gr[k].cpu().numpy().reshape(-1)) <= GRAD_RTOL, k
    assert float(gg["dscale"].abs().max()) == 0.0 and gg["dsh"].numel() == 0


@needs_ref
@pytest.mark.parametrize("deg", [0, 1, 2])
def test_lower_sh_degrees(deg):
    scene = synth.make_config("tiny")
    inp = helpers.raster_inputs(scene, view=0)
    inp["degree"] = deg
    cin = helpers.to_cuda(inp)
    cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
    ours = helpers.ours_forward(cin)
    gg = helpers.ours_backward(cin, ours, cots)
    f, st, gr = _run_ref(cin, cots)
    assert helpers.rel_err(ours["color"].cpu().numpy(), f["color"].cpu().numpy()) <= IMG_RTOL
    for k in ("dsh", "dmean3D"):
        assert helpers.rel_err(gg[k].cpu().numpy().reshape(-1), gr[k].cpu().numpy().reshape(-1)) <= GRAD_RTOL, k


def test_edge_cases():
    """Empty input, everything culled, a single Gaussian, image smaller than one tile."""
    from riggs_b200.diff_gaussian_rasterization import _C
    dev = "cuda"
    e = torch.empty(0)
    cam = synth.look_at_camera(torch.tensor([0.0, -3.0, 0.0]), 40, 24, 0.7).to(dev)
    import math
    tfx, tfy = math.tan(cam.FoVx / 2), math.tan(cam.FoVy / 2)
    bg = torch.tensor([0.5, 0.25, 0.125], device=dev)

    def run(means, **kw):
        P = means.shape[0]
        return _C.rasterize_gaussians(bg, means, kw.get("colors", torch.full((P, 3), 0.5, device=dev)),
                                      torch.full((P, 1), 0.8, device=dev), torch.full((P, 3), 0.05, device=dev),
                                      torch.tensor([[1.0, 0, 0, 0]], device=dev).repeat(P, 1), 1.0, e,
                                      cam.world_view_transform, cam.full_proj_transform, tfx, tfy, 24, 40, e, 0,
                                      cam.camera_center, False, False)

    # P == 0: zero images, nothing launched (dgr/rasterize_points.cu:83)
    R, color, depth, alpha, radii, *_ = run(torch.zeros(0, 3, device=dev))
    assert R == 0 and float(color.abs().max()) == 0.0 and radii.numel() == 0
    # all behind the camera: background everywhere, R == 0
    R, color, depth, alpha, radii, *_ = run(torch.tensor([[0.0, -10.0, 0.0], [0.0, -3.1, 0.0]], device=dev))
    assert R == 0 and int(radii.max()) == 0
    assert torch.allclose(color, bg[:, None, None].expand_as(color)) and float(alpha.max()) == 0.0
    # one Gaussian at the origin covers the centre pixel
    R, color, depth, alpha, radii, *_ = run(torch.zeros(1, 3, device=dev))
    assert R > 0 and int(radii[0]) > 0 and float(alpha[0, 12, 20]) > 0.5
    # wrong shape is rejected like the reference's AT_ERROR
    with pytest.raises(RuntimeError):
        run(torch.zeros(4, 2, device=dev))


@needs_ref
def test_mark_visible():
    scene = synth.make_config("tiny")
    cam = scene.cameras[0].to("cuda")
    from riggs_b200.diff_gaussian_rasterization import _C
    pts = (scene.xyz * 3.0).cuda().contiguous()
    ours = _C.mark_visible(pts, cam.world_view_transform, cam.full_proj_transform)
    ref = ref_cuda.mark_visible(pts, cam.world_view_transform.contiguous(), cam.full_proj_transform.contiguous())
    assert torch.equal(ours, ref)
    assert 0 < int(ours.sum()) <= pts.shape[0]


@pytest.mark.parametrize("P,kind", [(1, "rand"), (31, "ties"), (2048, "rand"), (2049, "ties"), (300000, "depth"),
                                    (300000, "ties"), (1000003, "rand"), (70000, "equal"), (50000, "onebyte"),
                                    (300000, "depth_all_visible")])
def test_depth_sort_is_the_stable_sort(P, kind):
    """csrc/radix_sort.cu against torch's stable sort: ascending key bits, equal keys keep their order (the order
    the reference's stable radix sort over id-ordered emission defines), for ragged sizes, heavy ties, one repeated
    key, depth-like floats with culled (+inf) entries, and more tiles than SMs."""
    from riggs_b200 import _lib
    g = torch.Generator().manual_seed(P)
    if kind == "rand":
        keys = torch.randint(0, 2**31 - 1, (P,), generator=g, dtype=torch.int64) * 2 + torch.randint(0, 2, (P,), generator=g)
    elif kind == "ties":
        keys = torch.randint(0, 50, (P,), generator=g, dtype=torch.int64) * 65537
    elif kind == "equal":
        keys = torch.full((P,), 0x40A00000, dtype=torch.int64)
    elif kind == "onebyte":  # three of the four digits are the same for every key: those passes are copies
        keys = torch.randint(0, 256, (P,), generator=g, dtype=torch.int64) * 256 + 0x40A00011
    elif kind == "depth_all_visible":  # depths inside [2, 8): the top byte is 0x40 for every key
        keys = (2.5 + 3.0 * torch.rand(P, generator=g)).view(torch.int32).to(torch.int64)
    else:
        d = 2.5 + 3.0 * torch.rand(P, generator=g)
        d[torch.rand(P, generator=g) < 0.1] = float("inf")
        keys = d.view(torch.int32).to(torch.int64)
    dev_keys = torch.from_numpy(keys.numpy().astype(np.uint32).view(np.int32)).cuda()
    out_k = torch.empty_like(dev_keys)
    out_i = torch.empty_like(dev_keys)
    L = _lib.lib()
    temp = torch.empty(int(L.rg_debug_depth_sort_bytes(P)), dtype=torch.uint8, device="cuda")
    for _ in range(2):  # the second call reuses the (dirty) temp buffer
        _lib.check(L.rg_debug_depth_sort(P, dev_keys.data_ptr(), out_k.data_ptr(), out_i.data_ptr(), temp.data_ptr(),
                                         _lib.stream_ptr()), "rg_debug_depth_sort")
        torch.cuda.synchronize()
        want_k, want_i = torch.sort(keys, stable=True)
        got_k = torch.from_numpy(out_k.cpu().numpy().view(np.uint32).astype(np.int64))
        assert torch.equal(got_k, want_k)
        assert torch.equal(out_i.cpu().to(torch.int64), want_i)
