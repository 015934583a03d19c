"""GPU tests of distCUDA2 and of the reference-facing Python surface (GaussianRasterizer / render())."""
import math

import numpy as np
import pytest
import torch

import helpers
from oracle import oracle_py as orc
from oracle import ref_cuda
from riggs_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P", [5, 1000, 20_000])
def test_dist_cuda2_vs_bruteforce_oracle(P):
    from riggs_b200.simple_knn._C import distCUDA2
    g = torch.Generator().manual_seed(P)
    pts = torch.randn(P, 3, generator=g)
    pts[: P // 4] = pts[: P // 4] * 0.01 + 2.0  # a dense cluster: very different box occupancy
    out = distCUDA2(pts.cuda()).cpu().numpy()
    ref = orc.dist2_knn3(pts.numpy())
    assert np.allclose(out, ref, rtol=1e-5, atol=1e-12)
    if ref_cuda.available():
        r2 = ref_cuda.dist_cuda2(pts.cuda().contiguous()).cpu().numpy()
        assert np.allclose(out, r2, rtol=1e-5, atol=1e-12)


def test_dist_cuda2_duplicates_and_empty():
    from riggs_b200.simple_knn._C import distCUDA2
    assert distCUDA2(torch.zeros(0, 3, device="cuda")).numel() == 0
    pts = torch.tensor([[0.0, 0, 0]] * 4 + [[1.0, 0, 0]], device="cuda")
    out = distCUDA2(pts).cpu().numpy()
    assert np.allclose(out[:4], 0.0) and np.isclose(out[4], 1.0)


class _Pipe:
    debug = False
    convert_SHs_python = False
    compute_cov3D_python = False


def test_render_signature_and_autograd_matches_kernel_level_calls():
    """render() (reference signature) -> autograd grads equal the kernel-level backward chained by hand."""
    from riggs_b200.gaussian_renderer import GaussianParams, render
    from riggs_b200 import skeleton_warp as sw
    scene = synth.make_config("tiny")
    pc = GaussianParams.from_scene(scene)
    cam = scene.cameras[0].to("cuda")
    warp = sw.SkeletonWarp(scene.joints, scene.parents, scene.node_radius_raw, use_skinning_weight_mlp=False,
                           use_template_offsets=False)
    pose = {k: v.cuda().requires_grad_(True) for k, v in scene.poses[0].items()}
    d = warp.deform_by_pose(pc.get_xyz, pose, pc.motion_mask)
    bg = torch.tensor([0.1, 0.2, 0.3], device="cuda")
    out = render(cam, pc, _Pipe(), bg, d["d_xyz"], d["d_rotation"], d["d_scaling"])
    assert set(out) == {"render", "viewspace_points", "visibility_filter", "radii", "depth", "alpha", "bg_color"}
    assert out["render"].shape == (3, cam.image_height, cam.image_width) and out["radii"].dtype == torch.int32
    cots = [t.cuda() for t in synth.seeded_cotangents(cam.image_height, cam.image_width)]
    loss = (out["render"] * cots[0]).sum() + (out["depth"] * cots[1]).sum() + (out["alpha"] * cots[2]).sum()
    loss.backward()
    assert out["viewspace_points"].grad is not None and out["viewspace_points"].grad.shape == (scene.N, 3)
    assert float(out["viewspace_points"].grad[:, 2].abs().max()) == 0.0
    for p in (pc._xyz, pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest, warp._node_radius):
        assert p.grad is not None and torch.isfinite(p.grad).all()
    assert pose["local_rotation"].grad is not None and float(pose["local_rotation"].grad.abs().max()) > 0
    # same numbers from the CPU oracles chained by torch autograd
    from oracle import skin_oracle as so
    inp = helpers.raster_inputs(scene, view=0, bg=(0.1, 0.2, 0.3))
    o = helpers.oracle_forward(inp)
    assert helpers.rel_err(out["render"].detach().cpu().numpy(), o["color"]) <= 1e-4
    go = helpers.oracle_backward(o, inp, cots := [c.cpu() for c in cots])
    # dL/d_xyz = dL/dmeans3D (means3D = _xyz + d_xyz with d_xyz computed from detached x)
    assert helpers.rel_err(pc._xyz.grad.cpu().numpy(), go["dmean3D"]) <= 2e-3


def test_rasterizer_argument_errors():
    from riggs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    cam = synth.look_at_camera(torch.tensor([0.0, -3.0, 0.0]), 32, 32, 0.7).to("cuda")
    rs = GaussianRasterizationSettings(32, 32, 0.36, 0.36, torch.zeros(3, device="cuda"), 1.0, cam.world_view_transform,
                                       cam.full_proj_transform, 0, cam.camera_center, False, False)
    r = GaussianRasterizer(rs)
    m = torch.zeros(3, 3, device="cuda")
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r(m, m, torch.ones(3, 1, device="cuda"), scales=torch.ones(3, 3, device="cuda"),
          rotations=torch.ones(3, 4, device="cuda"))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, torch.ones(3, 1, device="cuda"), colors_precomp=m)
    vis = r.markVisible(m)
    assert vis.dtype == torch.bool and vis.shape == (3,)


def test_large_config_properties():
    """BASELINE-sized input (C2: 200k Gaussians, 800x800): size-independent properties of the bit-exact stages
    -- keys sorted, stable on ties, ranges partition the list, offsets are the scan of tiles_touched --
    plus linearity of the backward in the cotangents."""
    scene = synth.make_config("C2")
    cin = helpers.to_cuda(helpers.raster_inputs(scene, view=0))
    st = helpers.ours_forward(cin)
    R = st["num_rendered"]
    assert R == int(st["tiles_touched"].astype(np.int64).sum())
    keys = st["keys"]
    assert np.all(keys[1:] >= keys[:-1])
    tie = keys[1:] == keys[:-1]
    assert np.all(st["point_list"][1:][tie] > st["point_list"][:-1][tie])  # stable: ties in id order
    # every Gaussian appears exactly tiles_touched times in the instance list
    assert np.array_equal(np.bincount(st["point_list"], minlength=len(st["tiles_touched"])), st["tiles_touched"])
    # the depth-ordered id list is a stable sort of the depths
    d = st["depths"].view(np.uint32)
    assert np.array_equal(st["sorted_ids"], np.argsort(d, kind="stable").astype(np.uint32))
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    rng = st["ranges"].astype(np.int64)
    lens = rng[:, 1] - rng[:, 0]
    assert lens.sum() == R and np.array_equal(np.bincount(tiles, minlength=rng.shape[0]), lens)
    nz = lens > 0
    assert np.all(tiles[rng[nz, 0]] == np.nonzero(nz)[0]) and np.all(tiles[rng[nz, 1] - 1] == np.nonzero(nz)[0])
    # depth bits of the key are the Gaussian's depth
    assert np.array_equal((keys & np.uint64(0xFFFFFFFF)).astype(np.uint32), st["depths"].view(np.uint32)[st["point_list"]])
    # image sanity: alpha in [0,1], n_contrib within the tile's list length
    a = st["alpha"].cpu().numpy()
    assert a.min() >= 0 and a.max() <= 1.0 + 1e-5
    # backward is linear in the cotangents
    c1 = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"], seed=1)]
    c2 = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"], seed=2)]
    g1 = helpers.ours_backward(cin, st, c1)
    g2 = helpers.ours_backward(cin, st, c2)
    g12 = helpers.ours_backward(cin, st, [x + 2 * y for x, y in zip(c1, c2)])
    for k in ("dmean3D", "dsh", "dscale", "drot", "dopacity"):
        lin = g1[k] + 2 * g2[k]
        assert helpers.rel_err(g12[k].cpu().numpy().reshape(-1), lin.cpu().numpy().reshape(-1)) <= 1e-4, k


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 800, 800), (3, 37, 53), (1, 1, 1), (3, 1, 5)])
def test_l1_loss_matches_torch(shape):
    """loss_utils.l1_loss: value within fp32 rounding of torch's mean |a-b|, gradient = sign(a-b)/n exactly,
    deterministic across calls, and usable twice in a row (the kernel re-arms its scratch)."""
    from riggs_b200.loss_utils import l1_loss
    g = torch.Generator().manual_seed(3)
    a = torch.rand(*shape, generator=g).cuda().requires_grad_(True)
    b = torch.rand(*shape, generator=g).cuda()
    with torch.no_grad():
        b.view(-1)[0] = a.view(-1)[0]  # a zero difference: gradient 0 there
    ref = (a.detach().double() - b.double()).abs().mean()
    losses = []
    for _ in range(2):
        a.grad = None
        loss = l1_loss(a, b)
        (loss * 2.0).backward()
        losses.append(float(loss.detach()))
        assert abs(float(loss.detach()) - float(ref)) <= 2e-6 * max(float(ref), 1e-6)
        want = torch.sign(a.detach() - b) * (1.0 / a.numel()) * 2.0
        assert torch.equal(a.grad, want.float())
    assert losses[0] == losses[1]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_p2p_kernel_on_one_device(world):
    """rg_allreduce_p2p with every 'rank' on this GPU (its peers' buffers are then ordinary local pointers): after each
    rank has reduced its slice, every copy holds the sum in rank order -- bit-identical on all ranks."""
    import ctypes as C
    from riggs_b200 import _lib
    n = 4 * 12345 + 4 * 3  # not a multiple of world * 4
    g = torch.Generator().manual_seed(world)
    bufs = [torch.randn(n, generator=g).cuda() for _ in range(world)]
    want = bufs[0].clone()
    for b in bufs[1:]:
        want = want + b
    arr = (C.c_void_p * world)(*[b.data_ptr() for b in bufs])
    L = _lib.lib()
    for r in range(world):
        _lib.check(L.rg_allreduce_p2p(arr, n, r, world, _lib.stream_ptr()), "rg_allreduce_p2p")
    torch.cuda.synchronize()
    for b in bufs:
        assert torch.equal(b, want)
    assert L.rg_allreduce_p2p(arr, n + 1, 0, world, _lib.stream_ptr()) < 0  # length must be a multiple of 4 floats
    assert L.rg_allreduce_p2p(arr, n, 0, 3, _lib.stream_ptr()) < 0


def test_l1_loss_and_grad_outside_autograd():
    from riggs_b200.loss_utils import l1_loss, l1_loss_and_grad
    g = torch.Generator().manual_seed(9)
    a = torch.rand(3, 64, 48, generator=g).cuda().requires_grad_(True)
    b = torch.rand(3, 64, 48, generator=g).cuda()
    loss, grad = l1_loss_and_grad(a, b)
    ref = l1_loss(a, b)
    ref.backward()
    assert float(loss) == float(ref.detach()) and torch.equal(grad, a.grad) and not loss.requires_grad


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_image_loss_vs_reference_golden(tag):
    """csrc/ssim.cu against the reference's own l1_loss / ssim / autograd outputs (tests/golden/loss_ssim.npz):
    loss, l1 and ssim within 1e-5, gradient within 1e-4 of its largest entry."""
    import os
    from riggs_b200.loss_utils import image_loss, image_loss_and_grad, ssim
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_ssim.npz"))
    img = torch.from_numpy(g[f"{tag}_img"]).float().cuda().requires_grad_(True)
    gt = torch.from_numpy(g[f"{tag}_gt"]).float().cuda()
    lam = float(g[f"{tag}_lambda"])
    loss, grad, l1, s = image_loss_and_grad(img, gt, lam)
    assert abs(float(loss) - float(g[f"{tag}_loss"])) <= 1e-5
    assert abs(float(l1) - float(g[f"{tag}_l1"])) <= 1e-5
    assert abs(float(s) - float(g[f"{tag}_ssim"])) <= 1e-5
    assert helpers.rel_err(grad.cpu().numpy(), g[f"{tag}_grad"]) <= 1e-4
    # the autograd wrappers
    l2 = image_loss(img, gt, lam)
    (l2 * 3.0).backward()
    assert float(l2.detach()) == float(loss) and torch.equal(img.grad, grad * 3.0)
    img.grad = None
    s2 = ssim(img, gt)
    s2.backward()
    assert abs(float(s2.detach()) - float(g[f"{tag}_ssim"])) <= 1e-5


@pytest.mark.parametrize("shape", [(3, 800, 800), (2, 3, 45, 31), (3, 1, 1), (3, 16, 17)])
def test_image_loss_vs_torch_oracle(shape):
    """Same kernels against oracle/loss_oracle.py (the reference's code restated) run in float64 on the same inputs,
    at the bench image size, for a batch, and for images smaller than the window."""
    from oracle import loss_oracle as lo
    from riggs_b200.loss_utils import image_loss_and_grad
    g = torch.Generator().manual_seed(11)
    gt = torch.rand(*shape, generator=g)
    img = (gt + 0.1 * torch.randn(*shape, generator=g)).clamp(0, 1)
    a = img.double().requires_grad_(True)
    want, wl1, ws = lo.image_loss(a, gt.double(), 0.2)
    want.backward()
    loss, grad, l1, s = image_loss_and_grad(img.cuda(), gt.cuda(), 0.2)
    want, wl1, ws = want.detach(), wl1.detach(), ws.detach()
    assert abs(float(loss) - float(want)) <= 1e-5 and abs(float(l1) - float(wl1)) <= 1e-5 and abs(float(s) - float(ws)) <= 1e-5
    assert helpers.rel_err(grad.cpu().numpy(), a.grad.numpy()) <= 1e-4


@pytest.mark.parametrize("N", [5000, 5001])  # 16-byte-word path and the scalar path
def test_fused_adam_matches_torch_adam(N):
    """optim.FusedAdam against torch.optim.Adam on the reference's group layout (six groups, own learning rates, eps
    1e-15, a learning-rate change between steps) over several steps."""
    from riggs_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(21)
    shapes = {"xyz": (N, 3), "f_dc": (N, 1, 3), "f_rest": (N, 15, 3), "opacity": (N, 1), "scaling": (N, 3),
              "rotation": (N, 4)}
    lrs = {"xyz": 1.6e-4 * 5, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 0.001 * 5, "rotation": 0.001}
    init = {k: torch.randn(*s, generator=g) for k, s in shapes.items()}
    a = {k: v.clone().cuda().requires_grad_(True) for k, v in init.items()}
    b = {k: v.clone().cuda().requires_grad_(True) for k, v in init.items()}
    mk = lambda d: [{"params": [d[k]], "lr": lrs[k], "name": k} for k in shapes]
    ours, ref = FusedAdam(mk(a), lr=0.0, eps=1e-15), torch.optim.Adam(mk(b), lr=0.0, eps=1e-15)
    for it in range(5):
        for k in shapes:
            gr = torch.randn(*shapes[k], generator=g).cuda() * (10.0 ** (it - 3))
            if k == "opacity":
                gr[::7] = 0.0  # zero gradients: the step is m / (sqrt(v) + 1e-15)
            a[k].grad, b[k].grad = gr.clone(), gr.clone()
        if it == 2:
            for opt in (ours, ref):
                for grp in opt.param_groups:
                    if grp["name"] == "xyz":
                        grp["lr"] = 3e-4
        ours.step()
        ref.step()
    torch.cuda.synchronize()
    for k in shapes:
        assert helpers.rel_err(a[k].detach().cpu().numpy(), b[k].detach().cpu().numpy()) <= 2e-6, k
        sa, sb = ours.state[a[k]], ref.state[b[k]]
        assert helpers.rel_err(sa["exp_avg"].cpu().numpy(), sb["exp_avg"].cpu().numpy()) <= 2e-6
        assert helpers.rel_err(sa["exp_avg_sq"].cpu().numpy(), sb["exp_avg_sq"].cpu().numpy()) <= 2e-6


def test_densification_stats_match_reference_expressions():
    from riggs_b200.optim import add_densification_stats
    g = torch.Generator().manual_seed(22)
    P = 10_001
    vsp = torch.zeros(P, 3).cuda().requires_grad_(True)
    vsp.grad = torch.randn(P, 3, generator=g).cuda()
    radii = (torch.randint(-2, 40, (P,), generator=g)).to(torch.int32).cuda()
    accum = torch.rand(P, 1, generator=g).cuda()
    denom = torch.rand(P, 1, generator=g).cuda().round()
    maxr = (torch.rand(P, generator=g) * 30).cuda()
    w_accum, w_denom, w_maxr = accum.clone(), denom.clone(), maxr.clone()
    vis = radii > 0
    # scene/gaussian_model.py:516-518 and train_rig.py:335
    w_accum[vis] += torch.norm(vsp.grad[vis, :2], dim=-1, keepdim=True)
    w_denom[vis] += 1
    w_maxr[vis] = torch.max(w_maxr[vis], radii[vis].float())
    add_densification_stats(vsp, radii, accum, denom, maxr)
    torch.cuda.synchronize()
    assert helpers.rel_err(accum.cpu().numpy(), w_accum.cpu().numpy()) <= 1e-6
    assert torch.equal(denom, w_denom) and torch.equal(maxr, w_maxr)


def test_prune_points_matches_boolean_indexing():
    """optim.prune_points against the reference's tensor[valid_points_mask] on parameters, Adam moments and statistics
    (bit-exact: pure data movement), including the all-kept, none-kept and single-row cases."""
    from riggs_b200.optim import FusedAdam, compact_rows, prune_points
    g = torch.Generator().manual_seed(31)
    P = 12_345
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    params = {k: torch.randn(*s, generator=g).cuda().requires_grad_(True) for k, s in shapes.items()}
    opt = FusedAdam([{"params": [params[k]], "lr": 1e-3, "name": k} for k in shapes], lr=0.0, eps=1e-15)
    for p in params.values():
        p.grad = torch.randn_like(p)
    opt.step()
    stats = {"xyz_gradient_accum": torch.rand(P, 1, generator=g).cuda(), "denom": torch.rand(P, 1, generator=g).cuda(),
             "max_radii2D": torch.rand(P, generator=g).cuda()}
    mask = (torch.rand(P, generator=g) < 0.3).cuda()
    valid = ~mask
    want_p = {k: v.detach()[valid].clone() for k, v in params.items()}
    want_m = {k: opt.state[v]["exp_avg"][valid].clone() for k, v in params.items()}
    want_v = {k: opt.state[v]["exp_avg_sq"][valid].clone() for k, v in params.items()}
    want_s = {k: v[valid].clone() for k, v in stats.items()}
    new_p, new_s = prune_points(opt, mask, stats)
    for grp in opt.param_groups:
        k = grp["name"]
        assert grp["params"][0] is new_p[k] and new_p[k].requires_grad and torch.equal(new_p[k].detach(), want_p[k])
        assert torch.equal(opt.state[new_p[k]]["exp_avg"], want_m[k]) and torch.equal(opt.state[new_p[k]]["exp_avg_sq"], want_v[k])
    for k in stats:
        assert torch.equal(new_s[k], want_s[k])
    # the optimizer keeps stepping on the pruned tensors
    for p in new_p.values():
        p.grad = torch.randn_like(p)
    opt.step()
    x = torch.arange(7 * 5, dtype=torch.float32).reshape(7, 5).cuda()
    for keep in (torch.ones(7, dtype=torch.bool), torch.zeros(7, dtype=torch.bool), torch.tensor([0, 0, 0, 1, 0, 0, 0]).bool()):
        (y,) = compact_rows(keep.cuda(), [x])
        assert torch.equal(y, x[keep.cuda()])


@pytest.mark.parametrize("isotropic", [False, True])
def test_densify_clone_split_prune_match_the_reference_sequence(isotropic):
    """optim.densify_and_clone / densify_and_split / densify_and_prune against the reference's own sequence restated with
    torch.cat and boolean indexing (scene/gaussian_model.py:393-514): same rows in the same order for every parameter
    (bit-exact for the copied ones, fp32 rounding for the children's sampled positions), zero Adam moments for new
    points, moments of surviving points carried along, statistics reset."""
    from riggs_b200.optim import FusedAdam, build_rotation, densify_and_clone, densify_and_prune, densify_and_split
    g = torch.Generator().manual_seed(77)
    P, extent, pd, thr, N = 8_001, 2.0, 0.01, 0.5, 2
    sc_cols = 1 if isotropic else 3
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, sc_cols),
              "rotation": (P, 4)}
    init = {k: torch.randn(*s, generator=g) for k, s in shapes.items()}
    init["scaling"] = torch.log(0.02 * torch.exp(0.7 * torch.randn(P, sc_cols, generator=g)))  # around percent_dense * extent
    params = {k: v.cuda().requires_grad_(True) for k, v in init.items()}
    opt = FusedAdam([{"params": [params[k]], "lr": 1e-3, "name": k} for k in shapes], lr=0.0, eps=1e-15)
    for p in params.values():
        p.grad = torch.randn_like(p)
    opt.step()
    stats = {"xyz_gradient_accum": torch.rand(P, 1, generator=g).cuda(), "denom": torch.ones(P, 1).cuda(),
             "max_radii2D": torch.rand(P, generator=g).cuda()}
    grads = stats["xyz_gradient_accum"] / stats["denom"]

    def act(s):
        return torch.exp(s.repeat(1, 3) if s.shape[1] == 1 else s)

    # ---- clone, the reference's way
    cur = {k: v.detach().clone() for k, v in params.items()}
    mom = {k: opt.state[v]["exp_avg"].clone() for k, v in params.items()}
    sel = (torch.norm(grads, dim=-1) >= thr) & (act(cur["scaling"]).max(1).values <= pd * extent)
    assert 0 < int(sel.sum()) < P
    want = {k: torch.cat([v, v[sel]]) for k, v in cur.items()}
    want_m = {k: torch.cat([v, torch.zeros_like(v[sel])]) for k, v in mom.items()}
    new, st = densify_and_clone(opt, stats, grads, thr, extent, pd)
    for k in shapes:
        assert torch.equal(new[k].detach(), want[k]), k
        assert torch.equal(opt.state[new[k]]["exp_avg"], want_m[k]), k
        assert opt.param_groups[list(shapes).index(k)]["params"][0] is new[k]
    n1 = want["xyz"].shape[0]
    assert st["xyz_gradient_accum"].shape == (n1, 1) and st["max_radii2D"].shape == (n1,) and float(st["denom"].abs().sum()) == 0
    # ---- split, the reference's way (grads padded with zeros for the clones)
    cur, mom = want, want_m
    padded = torch.zeros(n1, device="cuda")
    padded[:P] = grads.squeeze()
    sel = (padded >= thr) & (act(cur["scaling"]).max(1).values > pd * extent)
    ns = int(sel.sum())
    assert 0 < ns < n1
    z = torch.randn(ns * N, 3, generator=g).cuda()
    stds = act(cur["scaling"])[sel].repeat(N, 1)
    rots = build_rotation(cur["rotation"][sel]).repeat(N, 1, 1)
    child = {k: v[sel].repeat(N, *([1] * (v.dim() - 1))) for k, v in cur.items()}
    child["xyz"] = torch.bmm(rots, (z * stds).unsqueeze(-1)).squeeze(-1) + cur["xyz"][sel].repeat(N, 1)
    child["scaling"] = torch.log(stds / (0.8 * N))[:, :sc_cols]
    keep = ~sel
    want = {k: torch.cat([cur[k][keep], child[k]]) for k in cur}
    want_m = {k: torch.cat([mom[k][keep], torch.zeros_like(child[k])]) for k in cur}
    new, st = densify_and_split(opt, st, grads, thr, extent, pd, N=N, samples=z)
    for k in shapes:
        if k in ("xyz", "scaling"):
            assert torch.allclose(new[k].detach(), want[k], rtol=1e-6, atol=1e-6), k
            assert torch.equal(new[k].detach()[:int(keep.sum())], want[k][:int(keep.sum())]), k
        else:
            assert torch.equal(new[k].detach(), want[k]), k
        assert torch.equal(opt.state[new[k]]["exp_avg"], want_m[k]), k
    assert st["denom"].shape == (want["xyz"].shape[0], 1)
    # the optimizer keeps stepping on the regrouped tensors
    for p in new.values():
        p.grad = torch.randn_like(p)
    opt.step()
    # ---- the whole densify_and_prune: sizes follow the reference's arithmetic
    n2 = new["xyz"].shape[0]
    stats2 = {"xyz_gradient_accum": torch.rand(n2, 1, generator=g).cuda(), "denom": torch.ones(n2, 1).cuda(),
              "max_radii2D": torch.zeros(n2).cuda()}
    stats2["denom"][::7] = 0  # 0 / 0 -> nan -> 0 as in the reference
    stats2["xyz_gradient_accum"][::7] = 0
    cur = {k: v.detach().clone() for k, v in new.items()}
    gr = stats2["xyz_gradient_accum"] / stats2["denom"]
    gr[gr.isnan()] = 0
    big = act(cur["scaling"]).max(1).values > pd * extent
    hot = torch.norm(gr, dim=-1) >= thr
    n_clone, n_split = int((hot & ~big).sum()), int((hot & big).sum())
    op = torch.sigmoid(cur["opacity"]).squeeze()
    survivors = torch.cat([op[~(hot & big)], op[hot & ~big], op[hot & big].repeat(N)]) >= 0.3
    gen = torch.Generator(device="cuda").manual_seed(5)
    newp, news = densify_and_prune(opt, stats2, thr, 0.3, extent, None, pd, generator=gen)
    assert newp["xyz"].shape[0] == int(survivors.sum()) and n_clone > 0 and n_split > 0
    assert news["denom"].shape[0] == newp["xyz"].shape[0]
    assert bool((torch.sigmoid(newp["opacity"].detach()) >= 0.3).all())


@pytest.mark.parametrize("N,M,D,K", [(5000, 512, 3, 3), (3001, 1024, 3, 1), (1000, 1000, 3, 9), (777, 1500, 5, 4),
                                     (64, 5, 3, 8), (10, 2500, 8, 16)])
def test_knn_points_vs_bruteforce(N, M, D, K):
    """knn.knn_points against a float64 brute-force search: same neighbours in the same order, squared distances to
    fp32 rounding; M < K pads with zeros as pytorch3d does; self-query (the reference's node graph, :1061) finds itself."""
    from riggs_b200.knn import knn_points
    g = torch.Generator().manual_seed(N + M)
    p1 = torch.randn(1, N, D, generator=g).cuda()
    p2 = p1[:, :M].clone() if (N == M) else torch.randn(1, M, D, generator=g).cuda()
    out = knn_points(p1, p2, None, None, K=K, return_nn=True)
    d = ((p1[0].double()[:, None] - p2[0].double()[None]) ** 2).sum(-1)
    k = min(K, M)
    wd, wi = torch.topk(d, k, dim=1, largest=False, sorted=True)
    assert out.dists.shape == (1, N, K) and out.idx.dtype == torch.int64
    assert torch.equal(out.idx[0, :, :k], wi)
    assert helpers.rel_err(out.dists[0, :, :k].cpu().numpy(), wd.cpu().numpy()) <= 1e-6
    assert torch.equal(out.knn[0, :, :k], p2[0][wi])
    if K > M:
        assert float(out.dists[0, :, M:].abs().max()) == 0.0 and int(out.idx[0, :, M:].max()) == 0
    if N == M:
        assert torch.equal(out.idx[0, :, 0], torch.arange(N, device="cuda"))


def test_knn_points_gradient():
    from riggs_b200.knn import knn_points
    g = torch.Generator().manual_seed(5)
    p1 = torch.randn(1, 300, 3, generator=g).cuda().requires_grad_(True)
    p2 = torch.randn(1, 40, 3, generator=g).cuda().requires_grad_(True)
    out = knn_points(p1, p2, K=3)
    out.dists.sum().backward()
    q1, q2 = p1.detach().clone().requires_grad_(True), p2.detach().clone().requires_grad_(True)
    d = ((q1[0][:, None] - q2[0][None]) ** 2).sum(-1)
    torch.topk(d, 3, dim=1, largest=False).values.sum().backward()
    assert helpers.rel_err(p1.grad.cpu().numpy(), q1.grad.cpu().numpy()) <= 1e-5
    assert helpers.rel_err(p2.grad.cpu().numpy(), q2.grad.cpu().numpy()) <= 1e-5


def test_render_flow_matches_reference_rasterizer_on_the_same_flow():
    """render_flow (gaussian_renderer/__init__.py:154-269): the flow / motion-mask colours are formed in torch exactly as
    the reference does and splatted through colors_precomp; checked against the reference's CUDA rasterizer fed with the
    same per-Gaussian tensors, and the flow of an unmoved scene must be zero."""
    import math
    from oracle import ref_cuda
    from riggs_b200 import synth
    from riggs_b200.gaussian_renderer import GaussianParams, render_flow
    scene = synth.make_config("small", n_views=2, soft_mask=True)
    pc = GaussianParams.from_scene(scene)
    cam1, cam2 = scene.cameras[0].to("cuda"), scene.cameras[1].to("cuda")
    g = torch.Generator().manual_seed(2)
    d1 = (0.02 * torch.randn(scene.N, 3, generator=g)).cuda()
    d2 = (0.02 * torch.randn(scene.N, 3, generator=g)).cuda()
    out = render_flow(pc, cam1, cam2, d1, d2, 0.0, 0.0)
    assert out["render"].shape == (3, cam1.image_height, cam1.image_width)
    assert int(out["visibility_filter"].sum()) > 0
    # same camera, same deformation: no flow in the first two channels, the mask in the third
    still = render_flow(pc, cam1, None, d1, d1, 0.0, 0.0)
    assert float(still["render"][:2].detach().abs().max()) == 0.0 and float(still["render"][2].detach().max()) > 0.0
    if ref_cuda.available():
        with torch.no_grad():
            xyz = pc.get_xyz.detach()
            proj = lambda p, c: (lambda h: h[..., :3] / h[..., -1:])(torch.cat([p, torch.ones_like(p[..., :1])], -1) @ c.full_proj_transform)
            flow = proj(xyz + d2, cam2) - proj(xyz + d1, cam1)
            flow[..., -1:] = pc.motion_mask
            ref = ref_cuda.RefRasterizer()
            f = ref.forward(torch.zeros(3, device="cuda"), (xyz + d1).contiguous(), pc.get_opacity.detach().contiguous(),
                            cam1.world_view_transform.contiguous(), cam1.full_proj_transform.contiguous(),
                            cam1.camera_center.contiguous(), cam1.image_width, cam1.image_height, math.tan(cam1.FoVx * 0.5),
                            math.tan(cam1.FoVy * 0.5), colors_precomp=flow.contiguous(), scales=pc.get_scaling.detach().contiguous(),
                            rotations=pc.get_rotation.detach().contiguous(), degree=0)
        assert torch.equal(out["radii"], f["radii"])
        for k, kk in (("render", "color"), ("depth", "depth"), ("alpha", "alpha")):
            assert helpers.rel_err(out[k].detach().cpu().numpy(), f[kk].cpu().numpy()) <= 1e-5, k
    # the flow is differentiable w.r.t. the deformation that moves the target positions
    d2g = d2.clone().requires_grad_(True)
    render_flow(pc, cam1, cam2, d1, d2g, 0.0, 0.0)["render"][:2].sum().backward()
    assert d2g.grad is not None and torch.isfinite(d2g.grad).all() and float(d2g.grad.abs().max()) > 0


def test_fused_adam_keeps_a_step_count_per_parameter_like_torch():
    """torch.optim.Adam advances state['step'] only for parameters that have a gradient: a parameter that misses gradients
    on some iterations gets its own bias correction."""
    from riggs_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(3)
    a0, b0 = torch.randn(1000, 3, generator=g).cuda(), torch.randn(500, generator=g).cuda()
    pa, pb = a0.clone().requires_grad_(True), b0.clone().requires_grad_(True)
    qa, qb = a0.clone().requires_grad_(True), b0.clone().requires_grad_(True)
    ours = FusedAdam([{"params": [pa], "lr": 1e-2, "name": "a"}, {"params": [pb], "lr": 3e-3, "name": "b"}], lr=0.0, eps=1e-15)
    ref = torch.optim.Adam([{"params": [qa], "lr": 1e-2}, {"params": [qb], "lr": 3e-3}], lr=0.0, eps=1e-15)
    for it in range(6):
        ga, gb = torch.randn(1000, 3, generator=g).cuda(), torch.randn(500, generator=g).cuda()
        pa.grad, qa.grad = ga.clone(), ga.clone()
        if it % 3 != 1:  # b has no gradient on iterations 1 and 4
            pb.grad, qb.grad = gb.clone(), gb.clone()
        else:
            pb.grad = qb.grad = None
        ours.step()
        ref.step()
    torch.cuda.synchronize()
    assert ours.state[pb]["step"] == 4 and ours.state[pa]["step"] == 6
    assert helpers.rel_err(pa.detach().cpu().numpy(), qa.detach().cpu().numpy()) <= 1e-6
    assert helpers.rel_err(pb.detach().cpu().numpy(), qb.detach().cpu().numpy()) <= 1e-6
