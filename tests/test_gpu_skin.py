"""GPU parity of forward kinematics + skinning (CUDA kernels through the C ABI) against
  * the golden vectors captured from the REAL reference (tests/golden/skin_*.npz, made by
    tests/golden/make_skin_golden.py from skeleton_utils/skeleton_warp.py on the CPU), and
  * the torch restatement oracle/skin_oracle.py on larger seeded scenes, forward and backward.
Tolerances: 1e-5 relative on outputs, 1e-4 relative on gradients (fp32 sums in a different order)."""
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import skin_oracle as so
from riggs_b200 import synth

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
OUT_RTOL = 1e-5
GRAD_RTOL = 1e-4


def _load(name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLD, name + ".npz")).items()}


def _run_cuda(d, want_grad=True, blend_mode=0, K=-1):
    from riggs_b200 import skeleton_warp as sw
    dev = "cuda"
    f = lambda k: d[k].float().to(dev) if k in d and d[k] is not None else None
    q = f("local_rotation").requires_grad_(want_grad)
    g = f("global_trans").requires_grad_(want_grad)
    nr = f("node_radius_raw").requires_grad_(want_grad)
    mask = f("motion_mask").reshape(-1).requires_grad_(want_grad)
    s = f("mlp_weight")
    o = f("template_offsets")
    if s is not None:
        s.requires_grad_(want_grad)
    if o is not None:
        o.requires_grad_(want_grad)
    joints = f("joints")
    par = d["parents"].to(torch.int32).to(dev)
    G, node_rot, d_nodes = sw.fk(q, g, joints, par)
    if K > 0:
        d_xyz, d_rot, w, nn_idx, nn_d2 = sw.skin_topk(f("xyz"), joints, par, nr, G, node_rot, g, mask, o, K)
    else:
        d_xyz, d_rot, w = sw.skin(f("xyz"), joints, par, nr, G, node_rot, g, mask, s, o, blend_mode, True)
        nn_idx = nn_d2 = None
    out = dict(d_xyz=d_xyz, d_rotation=d_rot, d_nodes=d_nodes, nn_weight=w, G=G, node_rot=node_rot, nn_idx=nn_idx,
               nn_dist2=nn_d2)
    if want_grad:
        loss = (d_xyz * f("cot_xyz")).sum() + (d_rot * f("cot_rot")).sum()
        if "cot_nodes" in d:
            loss = loss + (d_nodes * f("cot_nodes")).sum()
        loss.backward()
        out.update(g_local_rotation=q.grad, g_global_trans=g.grad, g_node_radius=nr.grad, g_motion_mask=mask.grad,
                   g_mlp_weight=None if s is None else s.grad, g_template_offsets=None if o is None else o.grad)
    return out


@pytest.mark.parametrize("name", ["skin_f64_j7", "skin_f64_j24_plain", "skin_f32_j24", "skin_f32_j20_plain"])
def test_skin_vs_reference_golden(name):
    d = _load(name)
    out = _run_cuda(d)
    for k in ("d_xyz", "d_rotation", "d_nodes", "nn_weight"):
        assert helpers.rel_err(out[k].detach().cpu().numpy(), d[k].numpy()) <= OUT_RTOL, k
    for k in ("g_local_rotation", "g_global_trans", "g_node_radius", "g_motion_mask", "g_mlp_weight",
              "g_template_offsets"):
        if k in d:
            a = out[k].detach().cpu().numpy().reshape(-1)
            assert helpers.rel_err(a, d[k].numpy().reshape(-1)) <= GRAD_RTOL, k


def _topk_oracle_grads(d, nn_idx, K):
    """float64 oracle with the selection the kernel made, same cotangents."""
    dd = lambda k: d[k].double() if k in d else None
    q = dd("local_rotation").requires_grad_(True)
    g = dd("global_trans").requires_grad_(True)
    nr = dd("node_radius_raw").requires_grad_(True)
    mask = dd("motion_mask").requires_grad_(True)
    o = dd("template_offsets")
    if o is not None:
        o.requires_grad_(True)
    ref = so.deform_by_pose(dd("xyz"), dd("joints"), d["parents"], nr, q, g, mask, None, o, K=K, nn_idx=nn_idx)
    ((ref["d_xyz"] * dd("cot_xyz")).sum() + (ref["d_rotation"] * dd("cot_rot")).sum()).backward()
    return ref, dict(g_local_rotation=q.grad, g_global_trans=g.grad, g_node_radius=nr.grad, g_motion_mask=mask.grad,
                     g_template_offsets=None if o is None else o.grad)


@pytest.mark.parametrize("name", ["skin_f64_j24_k3", "skin_f32_j20_k4"])
def test_skin_topk_vs_reference_golden(name):
    """K nearest bones only (cal_nn_weight_skeleton with K > 0).  A point whose nearest point on two bones is their
    shared joint is exactly equidistant from both, and which of them torch.topk keeps is arbitrary: rows where the
    kernel selected the same SET of bones as the reference must give the reference's outputs; the gradients are
    checked over all rows against the float64 oracle (itself pinned to the reference in tests/test_oracle.py) run with
    the kernel's selection."""
    d = _load(name)
    K = d["nn_idx"].shape[1]
    out = _run_cuda(d, K=K)
    got_idx = out["nn_idx"].cpu()
    same = (torch.sort(got_idx, 1).values == torch.sort(d["nn_idx"], 1).values).all(1).numpy()
    assert same.mean() >= 0.75  # J = 20 random tree: every eighth point sits in a joint-nearest region with a tie at the K-th place
    for k in ("d_xyz", "d_rotation"):
        assert helpers.rel_err(out[k].detach().cpu().numpy()[same], d[k].numpy()[same]) <= OUT_RTOL, k
    assert helpers.rel_err(out["d_nodes"].detach().cpu().numpy(), d["d_nodes"].numpy()) <= OUT_RTOL
    ref, grads = _topk_oracle_grads(d, got_idx, K)
    for k in ("d_xyz", "d_rotation", "nn_weight", "nn_dist2"):
        assert helpers.rel_err(out[k].detach().cpu().numpy(), ref[k].detach().numpy(), floor=1e-7) <= OUT_RTOL, k
    for k, r in grads.items():
        if r is not None:
            a = out[k].detach().cpu().numpy().reshape(-1)
            assert helpers.rel_err(a, r.numpy().reshape(-1), floor=1e-6) <= GRAD_RTOL, k


def test_skin_topk_vs_oracle_and_module():
    """A larger cloud against the float64 oracle, through SkeletonWarp(K=3).deform_by_pose; K = J - 1 equals all bones."""
    from riggs_b200 import skeleton_warp as sw
    sc = synth.make_scene(N=40_000, J=24, width=64, height=64, n_views=1, seed=41, soft_mask=True, with_offsets=True)
    m = sw.SkeletonWarp(sc.joints, sc.parents, sc.node_radius_raw, use_skinning_weight_mlp=False,
                        use_template_offsets=False, K=3)
    pose = {k: v.cuda().requires_grad_(True) for k, v in sc.poses[0].items()}
    mask = sc.motion_mask.cuda().requires_grad_(True)
    out = m.deform_by_pose(sc.xyz.cuda(), pose, mask)
    gen = torch.Generator().manual_seed(42)
    ca, ch = torch.randn(sc.N, 3, generator=gen), torch.randn(sc.N, 4, generator=gen)
    ((out["d_xyz"] * ca.cuda()).sum() + (out["d_rotation"] * ch.cuda()).sum()).backward()
    dd = lambda t: t.double()
    q = dd(sc.poses[0]["local_rotation"]).requires_grad_(True)
    g = dd(sc.poses[0]["global_trans"]).requires_grad_(True)
    nr = dd(sc.node_radius_raw).requires_grad_(True)
    mk = dd(sc.motion_mask).requires_grad_(True)
    free = so.deform_by_pose(dd(sc.xyz), dd(sc.joints), sc.parents, nr, q, g, mk, None, None, K=3)
    # the kernel's distances are the K smallest (up to fp32 rounding), whatever it did with equidistant bones
    assert helpers.rel_err(out["nn_weight"].sum(1).cpu().numpy(), np.ones(sc.N)) <= 1e-6
    d2_sorted = torch.sort(free["nn_dist2"].detach(), 1).values.numpy()
    got_d2 = so.bone_dist2(dd(sc.xyz), dd(sc.joints), sc.parents).gather(1, out["nn_idx"].cpu() - 1)
    assert np.abs(torch.sort(got_d2, 1).values.numpy() - d2_sorted).max() <= 1e-6 * max(1.0, float(d2_sorted.max()))
    ref = so.deform_by_pose(dd(sc.xyz), dd(sc.joints), sc.parents, nr, q, g, mk, None, None, K=3, nn_idx=out["nn_idx"].cpu())
    ((ref["d_xyz"] * dd(ca)).sum() + (ref["d_rotation"] * dd(ch)).sum()).backward()
    for k in ("d_xyz", "d_rotation", "nn_weight"):
        assert helpers.rel_err(out[k].detach().cpu().numpy(), ref[k].detach().numpy()) <= OUT_RTOL, k
    for a, r in ((pose["local_rotation"].grad, q.grad), (pose["global_trans"].grad, g.grad),
                 (m._node_radius.grad, nr.grad), (mask.grad, mk.grad)):
        assert helpers.rel_err(a.cpu().numpy().reshape(-1), r.numpy().reshape(-1), floor=1e-6) <= GRAD_RTOL
    # all bones selected = the K = -1 kernel
    sc2 = synth.make_scene(N=3000, J=6, width=64, height=64, n_views=1, seed=43)
    f = lambda t: t.float().cuda()
    par = sc2.parents.to(torch.int32).cuda()
    G, node_rot, _ = sw.fk(f(sc2.poses[0]["local_rotation"]), f(sc2.poses[0]["global_trans"]), f(sc2.joints), par)
    a = sw.skin(f(sc2.xyz), f(sc2.joints), par, f(sc2.node_radius_raw), G, node_rot, f(sc2.poses[0]["global_trans"]))
    b = sw.skin_topk(f(sc2.xyz), f(sc2.joints), par, f(sc2.node_radius_raw), G, node_rot, f(sc2.poses[0]["global_trans"]), K=5)
    assert helpers.rel_err(b[0].cpu().numpy(), a[0].cpu().numpy()) <= OUT_RTOL
    assert helpers.rel_err(b[1].cpu().numpy(), a[1].cpu().numpy()) <= OUT_RTOL
    with pytest.raises(Exception):
        sw.skin_topk(f(sc2.xyz), f(sc2.joints), par, f(sc2.node_radius_raw), G, node_rot, f(sc2.poses[0]["global_trans"]), K=6)


@pytest.mark.parametrize("cfg", [dict(N=50_000, J=24, seed=31), dict(N=30_011, J=60, seed=32, soft_mask=True,
                                                                      with_mlp_weight=True, with_offsets=True),
                                 dict(N=1, J=2, seed=33), dict(N=257, J=3, seed=34, with_offsets=True)])
def test_skin_vs_torch_oracle(cfg):
    sc = synth.make_scene(width=64, height=64, n_views=1, **cfg)
    gen = torch.Generator().manual_seed(cfg["seed"] + 1)
    N, J = sc.N, sc.J
    d = dict(joints=sc.joints, parents=sc.parents, node_radius_raw=sc.node_radius_raw, xyz=sc.xyz,
             local_rotation=sc.poses[0]["local_rotation"], global_trans=sc.poses[0]["global_trans"],
             motion_mask=sc.motion_mask, cot_xyz=torch.randn(N, 3, generator=gen), cot_rot=torch.randn(N, 4, generator=gen),
             cot_nodes=torch.randn(J, 3, generator=gen))
    if sc.mlp_weight is not None:
        d["mlp_weight"] = sc.mlp_weight
    if sc.template_offsets is not None:
        d["template_offsets"] = sc.template_offsets
    out = _run_cuda(d)
    # float64 oracle
    dd = lambda k: d[k].double() if k in d else None
    q = dd("local_rotation").requires_grad_(True)
    g = dd("global_trans").requires_grad_(True)
    nr = dd("node_radius_raw").requires_grad_(True)
    mask = dd("motion_mask").requires_grad_(True)
    s, o = dd("mlp_weight"), dd("template_offsets")
    if s is not None:
        s.requires_grad_(True)
    if o is not None:
        o.requires_grad_(True)
    ref = so.deform_by_pose(dd("xyz"), dd("joints"), d["parents"], nr, q, g, mask, s, o)
    loss = (ref["d_xyz"] * dd("cot_xyz")).sum() + (ref["d_rotation"] * dd("cot_rot")).sum() + (ref["d_nodes"] * dd("cot_nodes")).sum()
    loss.backward()
    for k in ("d_xyz", "d_rotation", "d_nodes", "nn_weight"):
        assert helpers.rel_err(out[k].detach().cpu().numpy(), ref[k].detach().numpy()) <= OUT_RTOL, k
    G_ref = torch.cat([ref["G"][:, :3, :3].reshape(J, 9), ref["G"][:, :3, 3]], 1)
    assert helpers.rel_err(out["G"].detach().cpu().numpy(), G_ref.detach().numpy()) <= OUT_RTOL
    assert helpers.rel_err(out["node_rot"].cpu().numpy(), ref["node_rot"].numpy()) <= OUT_RTOL
    pairs = [("g_local_rotation", q.grad), ("g_global_trans", g.grad), ("g_node_radius", nr.grad),
             ("g_motion_mask", mask.grad), ("g_mlp_weight", None if s is None else s.grad),
             ("g_template_offsets", None if o is None else o.grad)]
    for k, r in pairs:
        if r is not None:
            # floor: with a single bone the normalised weight is 1 and several gradients are exactly zero
            assert helpers.rel_err(out[k].detach().cpu().numpy().reshape(-1), r.numpy().reshape(-1), floor=1e-6) <= GRAD_RTOL, k


def test_fk_batched_sequence():
    """F poses at once (novel-pose sequences) equal F single-pose calls and the oracle's chain."""
    from riggs_b200 import skeleton_warp as sw
    sc = synth.make_scene(N=16, J=60, width=32, height=32, n_views=1, n_poses=32, seed=5, pose_sequence=True)
    q = torch.stack([p["local_rotation"] for p in sc.poses]).cuda()
    t = torch.stack([p["global_trans"] for p in sc.poses]).cuda()
    par = sc.parents.to(torch.int32).cuda()
    G, node_rot, posed = sw.fk(q, t, sc.joints.cuda(), par)
    for f in (0, 7, 31):
        posed_ref, G_ref = so.forward_kinematics(so.quat_to_mat(q[f].cpu().double()), sc.joints.double(), sc.parents)
        G12 = torch.cat([G_ref[:, :3, :3].reshape(-1, 9), G_ref[:, :3, 3]], 1)
        assert helpers.rel_err(G[f].cpu().numpy(), G12.numpy()) <= OUT_RTOL
        assert helpers.rel_err(posed[f].cpu().numpy(), (posed_ref + t[f].cpu().double()).numpy()) <= OUT_RTOL
        assert helpers.rel_err(node_rot[f].cpu().numpy(), so.mat_to_quat(G_ref[:, :3, :3]).numpy()) <= OUT_RTOL


def test_dual_quaternion_blend_vs_reference_golden():
    """blend_mode=1 against utils/dual_quaternion.py (DQBlending), via the golden vectors."""
    from riggs_b200 import _lib
    from riggs_b200._lib import check, ptr, stream_ptr
    d = _load("dq_blend")
    K, N = d["q"].shape[0], d["w"].shape[0]
    # Feed the kernel a skeleton whose bone weights we control: weights come from mlp_weight with huge radii
    # (exp(-d2/2r2) == 1), so w_k = (s_k + 1e-7) / sum.
    J = K + 1
    dev = "cuda"
    joints = torch.zeros(J, 3)
    joints[1:, 0] = torch.arange(1, J) * 0.1
    parents = torch.tensor([-1] + [0] * K, dtype=torch.int32)
    G = torch.zeros(J, 12)
    G[1:, 9:] = d["t"]
    node_rot = torch.zeros(J, 4)
    node_rot[0, 0] = 1
    node_rot[1:] = d["q"]
    xyz = torch.randn(N, 3, generator=torch.Generator().manual_seed(1))
    d_xyz = torch.empty(N, 3, device=dev)
    d_rot = torch.empty(N, 4, device=dev)
    args = [t.to(dev).contiguous() for t in (xyz, joints, parents, torch.full((J,), 20.0), G, node_rot, torch.zeros(3))]
    w = d["w"].float().to(dev).contiguous()
    check(_lib.lib().rg_skin_forward(N, J, ptr(args[0]), ptr(args[1]), ptr(args[2]), ptr(args[3]), ptr(args[4]),
                                     ptr(args[5]), ptr(args[6]), None, ptr(w), None, 1, ptr(d_xyz), ptr(d_rot), None,
                                     stream_ptr()), "skin dq")
    R, T = d["R"], d["T"]
    expect = torch.einsum("nab,nb->na", R, xyz.double()) + T - xyz.double()
    assert helpers.rel_err(d_xyz.cpu().numpy(), expect.numpy()) <= 2e-5
    # blended unit quaternion: compare its rotation matrix (sign-free)
    Rq = so.quat_to_mat(d_rot.cpu().double())
    assert helpers.rel_err(Rq.numpy(), R.numpy()) <= 2e-5


@pytest.mark.parametrize("over", [dict(), dict(soft_mask=True, with_mlp_weight=True, with_offsets=True)])
def test_dual_quaternion_skinning_forward_backward_vs_float64_oracle(over):
    """blend_mode = 1 (rg_skin_forward / rg_skin_backward_blend) against the float64 oracle, whose dual-quaternion blend
    and its autograd gradients are pinned to the reference's DQBlending (tests/test_oracle.py).  Both sides start from the
    SAME bone table (the FK kernel's G and node_rot): the reference's formula standardises the sign of every bone's dual
    part by the sign of its real component, which is numerically zero for bones that barely move, so a bone table that
    differs in the last bit can flip a sign and change the blend -- a property of the reference's (unused) formula, not of
    either implementation."""
    from riggs_b200 import skeleton_warp as sw
    scene = synth.make_config("small", **over)
    g = torch.Generator().manual_seed(21)
    c_xyz = torch.randn(scene.N, 3, generator=g).double() / scene.N
    c_rot = torch.randn(scene.N, 4, generator=g).double() / scene.N
    joints = scene.joints.cuda()
    par = scene.parents.to(torch.int32).cuda()
    with torch.no_grad():
        G0, node_rot, _ = sw.fk(scene.poses[0]["local_rotation"].cuda(), scene.poses[0]["global_trans"].cuda(), joints, par)

    def leaves(dtype, dev):
        mk = lambda t: None if t is None else t.to(dev, dtype).clone().requires_grad_(True)
        return dict(nr=mk(scene.node_radius_raw), G=mk(G0.cpu()), g=mk(scene.poses[0]["global_trans"]),
                    mask=mk(scene.motion_mask), mlp=mk(scene.mlp_weight), off=mk(scene.template_offsets))

    L = leaves(torch.float64, "cpu")
    x64 = scene.xyz.double()
    w, _, _ = so.skin_weights(x64, scene.joints.double(), scene.parents, L["nr"], L["mlp"])
    q64, t64 = node_rot.cpu().double()[1:], L["G"][1:, 9:12]
    R, T = so.dq_blend(q64, t64, w)
    y = torch.einsum("nab,nb->na", R, x64) + T + L["g"] + (L["off"] if L["off"] is not None else 0.0)
    o_xyz = (y - x64) * L["mask"]
    o_rot = so.dq_blend_quat(q64, t64, w) * L["mask"]
    ((o_xyz * c_xyz).sum() + (o_rot * c_rot).sum()).backward()

    C = leaves(torch.float32, "cuda")
    d_xyz, d_rot, _ = sw.skin(scene.xyz.cuda(), joints, par, C["nr"], C["G"], node_rot, C["g"],
                              C["mask"].reshape(-1), C["mlp"], C["off"], 1, False)
    ((d_xyz * c_xyz.float().cuda()).sum() + (d_rot * c_rot.float().cuda()).sum()).backward()
    torch.cuda.synchronize()
    assert helpers.rel_err(d_xyz.detach().cpu().numpy(), o_xyz.detach().numpy()) <= 2e-5
    assert helpers.rel_err(d_rot.detach().cpu().numpy(), o_rot.detach().numpy()) <= 2e-5
    assert float(C["G"].grad[:, :9].abs().max()) == 0.0  # the rotation enters through the detached node_rot only
    for k in ("nr", "G", "g", "mask", "mlp", "off"):
        if L[k] is None:
            continue
        a, b = C[k].grad.cpu().numpy().reshape(-1), L[k].grad.numpy().reshape(-1)
        assert helpers.rel_err(a, b, floor=1e-12) <= 1e-4, k
        assert helpers.rel_l2(a, b) <= 1e-4, k
