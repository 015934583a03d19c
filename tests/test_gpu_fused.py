"""GPU parity of the fused entry points (rg_skinned_forward / rg_skinned_backward through render_skinned):
  * against the unfused path (SkeletonWarp.deform_by_pose + render(), i.e. the reference's call sequence on our
    kernels): images 1e-5, gradients 1e-4;
  * against the reference's CUDA rasterizer fed with the fused kernel's OWN means3D / scales / rotations /
    opacities (SURVEY.md "hard parts": torch's eager einsum order cannot be matched bit for bit, so the bit-exact
    gates of the fused entry are applied on identical per-Gaussian inputs): keys / order / ranges bit-exact."""
import math

import numpy as np
import pytest
import torch

import helpers
from oracle import ref_cuda
from riggs_b200 import synth

pytestmark = pytest.mark.gpu


class _Pipe:
    debug = False
    convert_SHs_python = False
    compute_cov3D_python = False


VARIANTS = [
    ("small", dict()),
    ("small", dict(soft_mask=True, with_mlp_weight=True, with_offsets=True)),
    ("small", dict(isotropic=True, soft_mask=True)),
    ("ragged", dict(with_offsets=True)),
]


def _setup(name, over):
    from riggs_b200 import skeleton_warp as sw
    from riggs_b200.gaussian_renderer import GaussianParams
    scene = synth.make_config(name, **over)
    pc = GaussianParams.from_scene(scene)
    warp = sw.SkeletonWarp(scene.joints, scene.parents, scene.node_radius_raw, use_skinning_weight_mlp=False,
                           use_template_offsets=False)
    cam = scene.cameras[0].to("cuda")
    return scene, pc, warp, cam


def _leafs(scene):
    pose = {k: v.cuda().clone().requires_grad_(True) for k, v in scene.poses[0].items()}
    mlp = scene.mlp_weight.cuda().clone().requires_grad_(True) if scene.mlp_weight is not None else None
    off = scene.template_offsets.cuda().clone().requires_grad_(True) if scene.template_offsets is not None else None
    return pose, mlp, off


def _loss(out, cots):
    return (out["render"] * cots[0]).sum() + (out["depth"] * cots[1]).sum() + (out["alpha"] * cots[2]).sum()


@pytest.mark.parametrize("name,over", VARIANTS)
def test_fused_matches_unfused(name, over):
    """The fused call against the reference's call sequence (deform_by_pose, activations, rasterize) on our unfused
    kernels.  torch's normalize differs from the in-kernel one in the last bit of some rotations, and the gradient
    of a 5-pixel splat amplifies one ulp of its rotation to ~1e-3 of the gradient, so the 1e-4 gate is applied the
    way the north star states it -- on IDENTICAL rasterizer inputs: the unfused rasterizer is run on the fused
    kernel's own activated tensors (bit-identical images expected), and its gradients are pushed through the
    torch/unfused-kernel chain.  The plain end-to-end comparison is kept as a looser sanity check."""
    from riggs_b200 import skeleton_warp as sw
    from riggs_b200.diff_gaussian_rasterization import _C
    from riggs_b200.fused import WorkspacePool, render_skinned
    from riggs_b200.gaussian_renderer import render
    scene, pc, warp, cam = _setup(name, over)
    P, W, H = scene.N, cam.image_width, cam.image_height
    bg = torch.tensor([0.1, 0.3, 0.2], device="cuda")
    cots = [t.cuda() for t in synth.seeded_cotangents(H, W)]
    cot_nodes = torch.randn(scene.J, 3, generator=torch.Generator().manual_seed(4)).cuda() * 1e-3
    params = [pc._xyz, pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest, warp._node_radius]
    iso = pc.use_isotropic_gs
    tfx, tfy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)

    def grads(pose, mlp, off, vsp):
        g = [p.grad.clone() for p in params] + [pose["local_rotation"].grad.clone(), pose["global_trans"].grad.clone(), vsp]
        opt = [mlp.grad.clone() if mlp is not None else None, off.grad.clone() if off is not None else None]
        for p in params:
            p.grad = None
        return g, opt

    # ---- fused
    pose, mlp, off = _leafs(scene)
    dbg = dict(scales=torch.empty(P, 3, device="cuda"), rotations=torch.empty(P, 4, device="cuda"),
               opacities=torch.empty(P, 1, device="cuda"))
    pool = WorkspacePool()
    out_f = render_skinned(cam, pc, warp, pose, _Pipe(), bg, mlp_weight=mlp, offsets=off, debug_outputs=dbg, pool=pool)
    (_loss(out_f, cots) + (out_f["d_nodes"] * cot_nodes).sum()).backward()
    g_f, g_f_opt = grads(pose, mlp, off, out_f["viewspace_points"].grad.clone())
    torch.cuda.synchronize()
    ws = pool.free[(P, W, H, str(pc._xyz.device))][0]
    ptrs = _C.state_ptrs(P, int(ws.status[0]), W, H, ws.geom, ws.binning, ws.image)
    means_f = helpers._slice(ws.geom, ptrs[7], 12 * P, torch.float32).reshape(P, 3).clone()

    # ---- unfused chain in torch / unfused kernels
    pose, mlp, off = _leafs(scene)
    joints = warp.nodes[:, :3]
    G, node_rot, d_nodes = sw.fk(pose["local_rotation"], pose["global_trans"], joints, warp.parents_i32)
    mask = pc._motion_mask.reshape(-1)
    d_xyz, d_rot, _ = sw.skin(pc.get_xyz, joints, warp.parents_i32, warp._node_radius, G, node_rot, pose["global_trans"],
                              mask, mlp, off)
    means3D = pc.get_xyz + d_xyz
    scales = pc.get_scaling
    rots = pc.get_rotation_bias(0.0 if iso else d_rot)
    opac = pc.get_opacity
    shs = pc.get_features
    assert torch.equal(means3D.detach(), means_f)          # same positions, bit for bit
    assert torch.equal(scales.detach(), dbg["scales"]) and torch.equal(opac.detach(), dbg["opacities"])
    assert float((rots.detach() - dbg["rotations"]).abs().max()) <= 2.5e-7
    # rasterizer fwd + bwd on the fused kernel's exact inputs
    e = torch.empty(0)
    R, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(
        bg, means_f, e, dbg["opacities"], dbg["scales"], dbg["rotations"], 1.0, e, cam.world_view_transform,
        cam.full_proj_transform, tfx, tfy, H, W, shs.detach().contiguous(), 3, cam.camera_center, False, False)
    assert R == int(ws.status[0]) and torch.equal(radii, out_f["radii"])
    # The two template instantiations of the preprocess kernel round the conic differently in the last bit for ~30% of
    # the Gaussians (measured: max 3.2e-7 relative; depth, means2D, rgb, radii are bit-identical).  Images agree to 2e-6.
    for a_, b_ in ((color, out_f["render"]), (depth, out_f["depth"]), (alpha, out_f["alpha"])):
        assert helpers.rel_err(b_.detach().cpu().numpy(), a_.cpu().numpy()) <= 2e-6
    # Gradients of this rasterizer are ill-conditioned in fp32: one ulp on a rotation moves the REFERENCE's own
    # gradients by 3e-4..5e-4 (measured, scratch/diag2.py).  So the backward kernels are compared on the SAME forward
    # state: the reference-shaped backward runs on the fused call's workspace (same buffer layout).
    # the fused workspace lays its binning buffer out for its CAPACITY (the survivor mask sits behind the list), so the
    # reference-shaped backward is told that number where it expects num_rendered
    from riggs_b200 import _lib
    cap = int(_lib.lib().rg_binning_capacity(ws.binning.numel()))
    gm2, _gc, gop, gm3, _gcov, gsh, gsc, grot = _C.rasterize_gaussians_backward(
        bg, means_f, out_f["radii"], e, dbg["scales"], dbg["rotations"], 1.0, e, cam.world_view_transform,
        cam.full_proj_transform, tfx, tfy, cots[0], cots[1], cots[2], shs.detach().contiguous(), 3, cam.camera_center,
        ws.geom, cap, ws.binning, ws.image, out_f["alpha"].detach(), False)
    torch.autograd.backward([means3D, scales, rots, opac, shs, d_nodes], [gm3, gsc, grot, gop, gsh, cot_nodes])
    g_u, g_u_opt = grads(pose, mlp, off, gm2)

    names = ["xyz", "scaling", "rotation", "opacity", "f_dc", "f_rest", "node_radius", "local_rot", "global_trans",
             "viewspace"]
    for n, a, b in zip(names, g_f, g_u):
        assert a.shape == b.shape, n
        floor = 1e-9
        if iso and n == "rotation":
            # an isotropic Gaussian's covariance does not depend on its (unit) rotation: the gradient is pure
            # cancellation noise around zero, so it is compared on the scale of the scaling gradient
            floor = float(g_u[1].abs().max())
        assert helpers.rel_err(a.cpu().numpy().reshape(-1), b.cpu().numpy().reshape(-1), floor=floor) <= 1e-4, n
    for n, a, b in zip(["mlp_weight", "offsets"], g_f_opt, g_u_opt):
        if b is not None:
            assert helpers.rel_err(a.cpu().numpy().reshape(-1), b.cpu().numpy().reshape(-1), floor=1e-9) <= 1e-4, n
    assert helpers.rel_err(out_f["d_nodes"].detach().cpu().numpy(), d_nodes.detach().cpu().numpy()) <= 1e-5

    # ---- looser end-to-end sanity: render() on torch-activated tensors (rotations may differ in the last bit)
    pose, mlp, off = _leafs(scene)
    G, node_rot, d_nodes = sw.fk(pose["local_rotation"], pose["global_trans"], joints, warp.parents_i32)
    d_xyz, d_rot, _ = sw.skin(pc.get_xyz, joints, warp.parents_i32, warp._node_radius, G, node_rot, pose["global_trans"],
                              mask, mlp, off)
    out_u = render(cam, pc, _Pipe(), bg, d_xyz, 0.0 if iso else d_rot, 0.0)
    (_loss(out_u, cots) + (d_nodes * cot_nodes).sum()).backward()
    g_e, _ = grads(pose, mlp, off, out_u["viewspace_points"].grad.clone())
    for k in ("render", "depth", "alpha"):
        assert helpers.rel_err(out_f[k].detach().cpu().numpy(), out_u[k].detach().cpu().numpy()) <= 1e-5, k
    for n, a, b in zip(names, g_f, g_e):
        floor = float(g_e[1].abs().max()) if (iso and n == "rotation") else 1e-9
        assert helpers.rel_err(a.cpu().numpy().reshape(-1), b.cpu().numpy().reshape(-1), floor=floor) <= 5e-3, n


@pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref/libref_riggs.so not built")
@pytest.mark.parametrize("name,over", VARIANTS[:3])
def test_fused_bit_exact_vs_reference_cuda_on_own_inputs(name, over):
    from riggs_b200.diff_gaussian_rasterization import _C
    from riggs_b200.fused import WorkspacePool, render_skinned
    scene, pc, warp, cam = _setup(name, over)
    P, W, H = scene.N, cam.image_width, cam.image_height
    bg = torch.tensor([0.0, 0.0, 0.0], device="cuda")
    pose = {k: v.cuda() for k, v in scene.poses[0].items()}
    dbg = dict(scales=torch.empty(P, 3, device="cuda"), rotations=torch.empty(P, 4, device="cuda"),
               opacities=torch.empty(P, 1, device="cuda"))
    pool = WorkspacePool()
    with torch.no_grad():
        out = render_skinned(cam, pc, warp, pose, _Pipe(), bg,
                             mlp_weight=scene.mlp_weight.cuda() if scene.mlp_weight is not None else None,
                             offsets=scene.template_offsets.cuda() if scene.template_offsets is not None else None,
                             debug_outputs=dbg, pool=pool)
    torch.cuda.synchronize()
    ws = pool.free[(P, W, H, str(pc._xyz.device))][0]
    R = int(ws.status[0])
    assert int(ws.status[1]) == 0
    ptrs = _C.state_ptrs(P, R, W, H, ws.geom, ws.binning, ws.image)
    means3D = helpers._slice(ws.geom, ptrs[7], 12 * P, torch.float32).reshape(P, 3).clone()
    shs = torch.cat([pc._features_dc, pc._features_rest], 1).detach().contiguous()
    ref = ref_cuda.RefRasterizer()
    f = ref.forward(bg, means3D, dbg["opacities"], cam.world_view_transform.contiguous(),
                    cam.full_proj_transform.contiguous(), cam.camera_center.contiguous(), W, H,
                    math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), shs=shs, scales=dbg["scales"],
                    rotations=dbg["rotations"])
    st = ref.state_arrays()
    assert R == f["num_rendered"]
    assert torch.equal(out["radii"], f["radii"])
    T = ((W + 15) // 16) * ((H + 15) // 16)
    point_list = helpers._slice(ws.binning, ptrs[8], 4 * R, torch.int32).cpu().numpy().view(np.uint32)
    keys = _C.debug_sorted_keys(P, R, W, H, ws.geom, ws.binning, ws.image).cpu().numpy().view(np.uint64)
    ranges = helpers._slice(ws.image, ptrs[10], 8 * T, torch.int32).cpu().numpy().view(np.uint32).reshape(T, 2)
    n_contrib = helpers._slice(ws.image, ptrs[9], 4 * W * H, torch.int32).cpu().numpy().view(np.uint32).reshape(H, W)
    assert np.array_equal(keys, st["keys"])
    assert np.array_equal(point_list, st["point_list"])
    assert np.array_equal(ranges, st["ranges"])
    assert np.array_equal(n_contrib, st["n_contrib"])
    for k, kk in (("render", "color"), ("depth", "depth"), ("alpha", "alpha")):
        assert helpers.rel_err(out[k].cpu().numpy(), f[kk].cpu().numpy()) <= 1e-5, k


def test_fused_capacity_growth_and_reuse():
    """A deliberately tiny first binning buffer overflows, is grown, and the frame is rendered again; later frames
    reuse the workspace without a host sync and give identical images."""
    from riggs_b200.fused import WorkspacePool, render_skinned
    scene, pc, warp, cam = _setup("tiny", {})
    bg = torch.zeros(3, device="cuda")
    pose = {k: v.cuda() for k, v in scene.poses[0].items()}
    pool = WorkspacePool(initial_bytes=4096)  # room for ~900 instances: the first frame must overflow
    key = (scene.N, cam.image_width, cam.image_height, str(pc._xyz.device))
    imgs = []
    with torch.no_grad():
        for _ in range(3):
            imgs.append(render_skinned(cam, pc, warp, pose, _Pipe(), bg, pool=pool)["render"].clone())
    torch.cuda.synchronize()
    assert torch.equal(imgs[0], imgs[1]) and torch.equal(imgs[1], imgs[2])
    assert len(pool.free[key]) == 1
    ws = pool.free[key][0]
    assert int(ws.status[1]) == 0 and int(ws.status[0]) > 0
    assert ws.binning.numel() >= 4 * int(ws.status[0])


def test_graphed_step_replays_the_eager_step():
    """HotPath.step captured as a CUDA graph gives, replay after replay and for new inputs copied into its static
    buffers, the loss and the gradients of the eager step."""
    from riggs_b200.hotpath import GraphedStep, HotPath
    scene = synth.make_config("small", n_views=2)
    hp = HotPath(scene, device="cuda")
    cams = [c.to("cuda") for c in scene.cameras[:2]]
    H, W = cams[0].image_height, cams[0].image_width
    g = torch.Generator().manual_seed(5)
    targets = [torch.rand(3, H, W, generator=g).cuda() for _ in range(2)]
    poses = [{k: v.cuda() for k, v in p.items()} for p in scene.poses[:2]]

    def eager(i):
        hp.zero_grad()
        ps = {k: v.clone().requires_grad_(True) for k, v in poses[i].items()}
        loss = hp.step(cams[i], ps, targets[i])
        return float(loss), {n: p.grad.clone() for n, p in hp.named_parameters()}, {k: v.grad.clone() for k, v in ps.items()}

    want = [eager(0), eager(1)]
    gs = GraphedStep(hp, cams[0], poses[0], targets[0])
    assert gs.launches_per_replay >= 8
    for n, i in enumerate((0, 1, 0, 1)):
        if n % 2 == 0:
            loss = gs.step(cam=cams[i], pose=poses[i], target=targets[i])
        else:  # inputs staged on the side stream (pinned host tensors), picked up by the next step()
            gs.prefetch(cam=cams[i], pose={k: v.cpu().pin_memory() for k, v in poses[i].items()},
                        target=targets[i].cpu().pin_memory())
            loss = gs.step()
        torch.cuda.synchronize()
        wl, wg, wp = want[i]
        assert abs(float(loss) - wl) <= 1e-6 * max(1.0, abs(wl))
        for n, p in hp.named_parameters():
            # float atomics in the blend backward make two runs of the same frame differ in the last bits
            assert helpers.rel_err(p.grad.cpu().numpy(), wg[n].cpu().numpy(), floor=1e-7) <= 1e-4, n
        for k, v in gs.pose_grads.items():
            assert helpers.rel_err(v.cpu().numpy(), wp[k].cpu().numpy(), floor=1e-7) <= 1e-4, k


def test_pipelined_steps_feed_from_host_and_read_losses_late():
    """PipelinedSteps: host inputs fed one step ahead into the idle slot, losses read one step late -- every step
    gives the loss and the gradients of the eager step on the same inputs."""
    from riggs_b200.hotpath import HotPath, PipelinedSteps
    scene = synth.make_config("small", n_views=2)
    hp = HotPath(scene, device="cuda")
    cams = [c.to("cuda") for c in scene.cameras[:2]]
    H, W = cams[0].image_height, cams[0].image_width
    g = torch.Generator().manual_seed(6)
    targets = [torch.rand(3, H, W, generator=g).cuda() for _ in range(2)]
    poses = [{k: v.cuda() for k, v in p.items()} for p in scene.poses[:2]]

    def eager(i):
        hp.zero_grad()
        ps = {k: v.clone().requires_grad_(True) for k, v in poses[i].items()}
        loss = hp.step(cams[i], ps, targets[i])
        return float(loss), {n: p.grad.clone() for n, p in hp.named_parameters()}, {k: v.grad.clone() for k, v in ps.items()}

    want = [eager(0), eager(1)]
    host = [(cams[i], {k: v.cpu().pin_memory() for k, v in poses[i].items()}, targets[i].cpu().pin_memory())
            for i in range(2)]
    order = (0, 1, 1, 0, 1)
    pipe = PipelinedSteps(hp, cams[0], poses[0], targets[0])
    pipe.feed(*host[order[0]])
    losses = {}
    for n, i in enumerate(order):
        assert pipe.launch() == n
        if n + 1 < len(order):
            pipe.feed(*host[order[n + 1]])
        if n > 0:
            losses[n - 1] = pipe.result(n - 1)
        torch.cuda.synchronize()
        wl, wg, wp = want[i]
        for name, p in hp.named_parameters():
            assert helpers.rel_err(p.grad.cpu().numpy(), wg[name].cpu().numpy(), floor=1e-7) <= 1e-4, name
        for k, v in pipe.pose_grads.items():
            assert helpers.rel_err(v.cpu().numpy(), wp[k].cpu().numpy(), floor=1e-7) <= 1e-4, k
    losses[len(order) - 1] = pipe.result(len(order) - 1)
    for n, i in enumerate(order):
        assert abs(losses[n] - want[i][0]) <= 1e-6 * max(1.0, abs(want[i][0])), n
