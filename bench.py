#!/usr/bin/env python
"""bench.py -- skin + splat-rasterize forward/backward iterations per second (BASELINE.json's metric).

    python bench.py --gpus N --steps K --warmup W            # our sm_100a path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port), host cores

A step = one pass of the hot path over one view: pose -> FK -> skin every Gaussian -> render 800x800 ->
L1 loss -> backward to all Gaussian parameters and the skeleton inputs (train_rig.py:535-554 minus SSIM).
Workload at every N: config C0 of SURVEY.md section 8(d) = the configuration the metric is quoted on
(300k Gaussians, 24-joint SMPL-like skeleton, 800x800), one view per GPU per step (weak scaling); with N > 1
the Gaussian-parameter gradients are summed across ranks by NCCL all-reduce inside the timed region.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "deform+raster fwd+bwd iters/s"
UNIT = "iters/s"
WORKLOAD = "C0"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--isotropic", action="store_true")
    ap.add_argument("--lambda-dssim", type=float, default=0.0,
                    help="> 0: the step's loss is the trainer's (1-l)*L1 + l*(1-SSIM) (fused kernels) instead of L1")
    ap.add_argument("--no-graph", action="store_true", help="drive the fused step launch by launch instead of replaying a CUDA graph")
    ap.add_argument("--unfused", action="store_true", help="reference call sequence (deform_by_pose + render) on our kernels")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------
# clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1] or [r for (_, r) in self.rows[-3:]]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
# algorithmic bytes (restated in DESIGN.md)
def algorithmic_bytes(P, R, HW, in_P=236, out_P=248, passes=6):
    per = {
        # fused path
        "skin_preprocess": P * (in_P + 48 + 12),
        "gauss_skin_backward": P * (in_P + 40 + 12 + out_P),
        # unfused path
        "preprocess": P * (in_P + 48),
        "gauss_backward": P * (in_P + 40 + out_P),
        "skin_forward": P * (12 + 28),
        "skin_backward": P * (12 + 28),
        # binning: depth sort of P (key,id) pairs, 4 digit passes; counts matrix; one 4-byte write per instance
        "depth_sort": P * 16 * 4,
        "tile_count": P * 20,
        "tile_offsets": 0,
        "tile_scatter": P * 20 + R * 4,
        "blend_forward": R * 44 + HW * 24,
        "blend_backward": HW * 28 + R * (44 + 40),
    }
    total = P * (2 * in_P + 336) + R * (148 + 24 * passes) + HW * 52
    return per, total


def make_scene(workload, isotropic=False):
    from riggs_b200 import synth
    return synth.make_config(workload, n_views=8, isotropic=isotropic)


# ---------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from riggs_b200 import _lib
    from riggs_b200.hotpath import HotPath

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        os.environ.setdefault("NCCL_DEBUG", "WARN")  # keep NCCL banner lines off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=120))
    scene = make_scene(args.workload, args.isotropic)
    hp = HotPath(scene, device=dev, fused=not args.unfused, lambda_dssim=args.lambda_dssim)
    cam = scene.cameras[rank % len(scene.cameras)].to(dev)
    H, W = cam.image_height, cam.image_width
    g = torch.Generator().manual_seed(1000 + rank)
    target_host = torch.rand(3, H, W, generator=g).pin_memory()
    target = target_host.to(dev)
    pose_host = {k: v.clone().pin_memory() for k, v in scene.poses[rank % len(scene.poses)].items()}
    pose = {k: v.to(dev).requires_grad_(True) for k, v in pose_host.items()}
    params = hp.parameters()
    n_grad = sum(p.numel() for p in params)
    if world > 1:
        # data parallel over views: the backward kernels write into one flat fp32 buffer, all-reduced in place by NCCL
        hp.enable_flat_grads()

    def reduce_grads():
        if world > 1:
            hp.reduce_grads()

    def eager_step(tgt, ps):
        hp.zero_grad()
        for v in ps.values():
            v.grad = None
        loss = hp.step(cam, ps, tgt)
        reduce_grads()
        return loss

    # The fused step is captured once as a CUDA graph (riggs_b200.hotpath.GraphedStep) and replayed; the gradient
    # all-reduce of the data-parallel run stays an eager NCCL call after each replay.
    use_graph = hp.fused and not args.no_graph
    gs = None
    if use_graph:
        from riggs_b200.hotpath import GraphedStep
        gs = GraphedStep(hp, cam, pose, target, warmup=3)

    def one_step(tgt, ps):
        if gs is None:
            return eager_step(tgt, ps)
        loss = gs.step()  # target / pose / camera already sit in the graph's static device buffers
        reduce_grads()
        return loss

    L = _lib.lib()
    # nvidia-smi takes a driver lock while it starts: start it before the warm-up, not inside the timed region
    clocks = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        one_step(target, pose)
    torch.cuda.synchronize()
    if clocks is not None:
        t_wait = time.time()
        while not clocks.rows and time.time() - t_wait < 5.0:
            time.sleep(0.05)
    for _ in range(3):  # every rank runs the same number of steps (each one contains a collective when N > 1)
        one_step(target, pose)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()

    # ---- timed region 1: inputs resident in HBM -------------------------------------------------------
    launches0 = L.rg_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        one_step(target, pose)
    e1.record()
    torch.cuda.synchronize()
    t1 = time.time()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    # kernels of this library inside the timed region (a replayed graph launches what its capture recorded)
    launches = gs.launches_per_replay * args.steps if gs is not None else L.rg_launch_count() - launches0
    clk = clocks.stop(t0, t1) if clocks else None

    # ---- per-kernel durations: a separate eager pass with every launch bracketed by CUDA events on its stream ----
    L.rg_profile_reset()
    L.rg_profile_enable(1)
    for _ in range(args.steps):
        eager_step(target, pose)
    torch.cuda.synchronize()
    L.rg_profile_enable(0)
    prof = _lib.profile_read()
    L.rg_profile_reset()
    if world > 1:
        dist.barrier()

    # ---- timed region 2: end to end through the public API, host buffers ------------------------------
    cam_host = [cam.world_view_transform.cpu().pin_memory(), cam.full_proj_transform.cpu().pin_memory(),
                cam.camera_center.cpu().pin_memory()]
    import copy
    cam2 = copy.copy(cam)
    class _HostCam:  # pinned host copies of the camera tensors
        world_view_transform, full_proj_transform, camera_center = cam_host

    pipe = None
    if gs is not None:  # two captured graphs alternate so that step i+1's inputs land while step i runs (set-up, untimed)
        from riggs_b200.hotpath import PipelinedSteps
        pipe = PipelinedSteps(hp, cam, pose, target, warmup=3, share_workspace=os.environ.get("RG_PIPE_SHARE", "1") != "0")
        for _ in range(2):
            pipe.feed(cam=_HostCam, pose=pose_host, target=target_host)
            pipe.result(pipe.launch())
            reduce_grads()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h2d = target_host.numel() * 4 + sum(v.numel() * 4 for v in pose_host.values()) + sum(c.numel() * 4 for c in cam_host)
    torch.cuda.synchronize()
    f0.record()
    last = 0.0
    if pipe is not None:  # input pipeline: step i+1's host -> device copies run while step i computes
        pipe.feed(cam=_HostCam, pose=pose_host, target=target_host)
    for it in range(args.steps):
        if pipe is not None:
            pipe.launch()  # waits for this step's inputs, replays, queues the read-back of the step's loss
            if it + 1 < args.steps:
                pipe.feed(cam=_HostCam, pose=pose_host, target=target_host)
            reduce_grads()
            if it > 0:
                last = pipe.result(it - 1)  # host read of the previous step's loss while this step runs
            continue
        tgt = target_host.to(dev, non_blocking=True)
        ps = {k: v.to(dev, non_blocking=True).requires_grad_(True) for k, v in pose_host.items()}
        cam2.world_view_transform = cam_host[0].to(dev, non_blocking=True)
        cam2.full_proj_transform = cam_host[1].to(dev, non_blocking=True)
        cam2.camera_center = cam_host[2].to(dev, non_blocking=True)
        hp.zero_grad()
        loss = hp.step(cam2, ps, tgt)
        reduce_grads()
        last = float(loss.cpu())  # device -> host read of the step's result
    if pipe is not None:
        last = pipe.result(args.steps - 1)
    f1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_e2e = f0.elapsed_time(f1)

    # max over ranks
    if world > 1:
        t = torch.tensor([ms, ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # workload facts for the roofline
    with torch.no_grad():
        out = hp.forward(cam, pose)
    P = scene.N
    vis = int((out["radii"] > 0).sum())
    from riggs_b200.diff_gaussian_rasterization import _C  # noqa: F401
    R = int(getattr(hp, "_last_R", 0))
    # num_rendered is returned by the C ABI; fetch it once more through the binding
    from riggs_b200 import gaussian_renderer as gr  # noqa: F401
    R = _count_rendered(hp, cam, pose)
    in_P = 236 - (8 if args.isotropic else 0)
    per, total_bytes = algorithmic_bytes(P, R, H * W, in_P=in_P)
    step_ms = ms / args.steps
    value = world * args.steps / (ms / 1e3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured" if "hbm_gbs" in peaks else "fallback"
    kern = {k: {"ms_per_launch": v[0] / max(v[1], 1), "launches": v[1], "ms_per_step": v[0] / args.steps}
            for k, v in prof.items()}
    dom = max(kern, key=lambda k: kern[k]["ms_per_step"]) if kern else None
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the one `ncu --set full` capture of this workload
    # (profiles/r1_t_full_raw.csv, C0 on one B200); other workloads have no capture -> null
    ncu_traffic = {"blend_backward": 17084416, "blend_forward": 14722304, "gauss_skin_backward": 122528256,
                   "skin_preprocess": 78285056} if args.workload == "C0" and not args.isotropic else {}
    roof = None
    if dom:
        b = per.get(dom, 0)
        achieved = b / (kern[dom]["ms_per_launch"] * 1e-3) / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": ncu_traffic.get(dom), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": b, "ms_per_launch": round(kern[dom]["ms_per_launch"], 4)}
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(step_ms, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {P} Gaussians, {scene.J}-joint skeleton, {W}x{H}, 1 view/GPU/step, "
                               f"fwd+bwd", "P": P, "visible": vis, "R": R, "HW": H * W,
                   "l2": "inputs+state per step (~%d MB) exceed no L2 flush; working set streams through HBM" % (total_bytes >> 20),
                   "launch": "cuda graph replay" if gs is not None else "eager",
                   "loss": "L1" if args.lambda_dssim <= 0 else f"(1-{args.lambda_dssim})*L1 + {args.lambda_dssim}*(1-SSIM)",
                   "parallelism": (f"dp{world} over views, " + (("own all-reduce kernel over symmetric memory (" + ("rg_allreduce_multimem, NVLS" if getattr(hp.flat, "_mc_ptr", 0)
                                                                                       else "rg_allreduce_p2p, NVLink peer loads/stores") + ")")
                                   if getattr(hp.flat, "hdl", None) is not None else "NCCL all-reduce") +
                                   f" of {n_grad} grad floats") if world > 1 else "single GPU"},
        "e2e": {"value": round(world * args.steps / (ms_e2e / 1e3), 3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "last_loss": last},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": roof,
        "step_hbm": {"algorithmic_bytes_per_step": total_bytes, "achieved_gbs": round(total_bytes / (step_ms * 1e-3) / 1e9, 2),
                     "frac_of_peak": round(total_bytes / (step_ms * 1e-3) / 1e9 / peak, 4), "peak": peak},
        "kernels": {k: {kk: round(vv, 4) if isinstance(vv, float) else vv for kk, vv in v.items()} for k, v in kern.items()},
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(scene, args)
    if world == 1 and not args.no_ref_cuda:
        rc = ref_cuda_baseline(scene, cam, pose, target, steps=min(args.steps, 10))
        if rc:
            line["reference_cuda_b200"] = rc
    if world == 1:
        line["next_rows"] = next_rows(hp, out["render"].detach(), target)
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


def next_rows(hp, render, target, iters=20):
    """Device time of the SURVEY 8(f) rows built so far, outside the headline step: the trainer's L1 + SSIM image loss with
    its gradient (N1) and the Adam step over every Gaussian parameter group (N2), CUDA events on the launching stream."""
    import torch
    from riggs_b200.loss_utils import image_loss_and_grad
    from riggs_b200.optim import FusedAdam
    names = ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
    ps = hp.parameters()[:6]
    clones = [p.detach().clone().requires_grad_(True) for p in ps]
    for c in clones:
        c.grad = torch.randn_like(c) * 1e-3
    opt = FusedAdam([{"params": [c], "lr": 1e-4, "name": n} for c, n in zip(clones, names)], lr=0.0, eps=1e-15)

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    n = sum(c.numel() for c in clones)
    ms_adam = timed(opt.step)
    ms_loss = timed(lambda: image_loss_and_grad(render, target, 0.2))
    return {"image_loss_l1_ssim_ms": round(ms_loss, 4), "adam_step_ms": round(ms_adam, 4), "adam_params": n,
            "adam_gbs": round(n * 28 / (ms_adam * 1e-3) / 1e9, 1)}


def _count_rendered(hp, cam, pose):
    """num_rendered of the bench view (through the reference-shaped binding)."""
    import math
    import torch
    from riggs_b200 import skeleton_warp as sw
    from riggs_b200.diff_gaussian_rasterization import _C
    with torch.no_grad():
        joints = hp.warp.nodes[:, :3]
        G, node_rot, _ = sw.fk(pose["local_rotation"], pose["global_trans"], joints, hp.warp.parents_i32)
        d_xyz, d_rot, _ = sw.skin(hp.pc.get_xyz, joints, hp.warp.parents_i32, hp.warp._node_radius, G, node_rot,
                                  pose["global_trans"])
        pc = hp.pc
        e = torch.empty(0)
        R, *_ = _C.rasterize_gaussians(hp.bg, pc.get_xyz + d_xyz, e, pc.get_opacity, pc.get_scaling,
                                       pc.get_rotation_bias(0.0 if hp.isotropic else d_rot), 1.0, e,
                                       cam.world_view_transform, cam.full_proj_transform,
                                       math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), cam.image_height,
                                       cam.image_width, pc.get_features, 3, cam.camera_center, False, False)
    return int(R)


# ---------------------------------------------------------------------------------------------------------
# CPU path: the oracle port (skin_oracle.py in torch on all host threads + raster_oracle.c with OpenMP)
def cpu_step(scene, view, target):
    """One full iteration on the host cores; returns the loss."""
    import numpy as np
    import torch
    from oracle import oracle_py as orc
    from oracle import skin_oracle as so
    import math
    cam = scene.cameras[view]
    p = scene.poses[view]
    leaf = lambda t: t.clone().requires_grad_(True)
    xyz, sc_raw, rot_raw, op_raw = leaf(scene.xyz), leaf(scene.scaling_raw), leaf(scene.rotation_raw), leaf(scene.opacity_raw)
    fdc, frest = leaf(scene.features_dc), leaf(scene.features_rest)
    q, g, nr = leaf(p["local_rotation"]), leaf(p["global_trans"]), leaf(scene.node_radius_raw)
    d = so.deform_by_pose(xyz, scene.joints, scene.parents, nr, q, g, scene.motion_mask)
    means3D, scales, rots, opac = so.activate(xyz, sc_raw, rot_raw, op_raw, d["d_xyz"], d["d_rotation"])
    shs = torch.cat([fdc, frest], 1)
    W, H = cam.image_width, cam.image_height
    tfx, tfy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
    view_m, proj_m, campos = cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), cam.camera_center.numpy()
    bg = np.zeros(3, np.float32)
    kw = dict(shs=shs.detach().numpy(), scales=scales.detach().numpy(), rotations=rots.detach().numpy())
    st = orc.forward(means3D.detach().numpy(), opac.detach().numpy(), view_m, proj_m, campos, W, H, tfx, tfy, bg, **kw)
    img = torch.from_numpy(st["color"])
    loss = (img - target).abs().mean()
    gimg = (torch.sign(img - target) / img.numel()).numpy()
    zeros = np.zeros((1, H, W), np.float32)
    gr = orc.backward(st, means3D.detach().numpy(), view_m, proj_m, campos, W, H, tfx, tfy, bg, gimg, zeros, zeros, **kw)
    torch.autograd.backward([means3D, scales, rots, opac, shs],
                            [torch.from_numpy(gr["dmean3D"]), torch.from_numpy(gr["dscale"]), torch.from_numpy(gr["drot"]),
                             torch.from_numpy(gr["dopacity"]).reshape(-1, 1), torch.from_numpy(gr["dsh"])])
    return float(loss)


def cpu_baseline(scene, args, budget_s=12.0):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    cam = scene.cameras[0]
    target = torch.rand(3, cam.image_height, cam.image_width, generator=torch.Generator().manual_seed(1000))
    cpu_step(scene, 0, target)  # warm-up (also builds the oracle library)
    n, t0 = 0, time.time()
    while True:
        cpu_step(scene, 0, target)
        n += 1
        if time.time() - t0 > budget_s or n >= 40:  # a bounded sample: ~10-15 s of host work
            break
    dt = (time.time() - t0) / n
    return {"value": round(1.0 / dt, 4), "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} full {args.workload} iterations (skin_oracle.py torch-CPU + raster_oracle.c OpenMP), {dt:.2f} s each"}


def run_reference(args):
    """The reference's CPU implementation of the path, as its oracle port, on all host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene = make_scene(args.workload, args.isotropic)
    cam = scene.cameras[0]
    target = torch.rand(3, cam.image_height, cam.image_width, generator=torch.Generator().manual_seed(1000))
    for _ in range(min(args.warmup, 1)):
        cpu_step(scene, 0, target)
    t0 = time.time()
    steps = 0
    for _ in range(args.steps):
        cpu_step(scene, 0, target)
        steps += 1
        if time.time() - t0 > 150:  # keep the whole run within a few minutes
            break
    dt = (time.time() - t0) / steps
    v = round(1.0 / dt, 4)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": min(args.warmup, 1), "ms_per_step": round(dt * 1e3, 2), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {scene.N} Gaussians, {scene.J}-joint skeleton, "
                                   f"{cam.image_width}x{cam.image_height}, 1 view/step, fwd+bwd on host cores"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{steps} full iterations, oracle port (torch-CPU skinning + OpenMP C rasterizer)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(line)


def ref_cuda_baseline(scene, cam, pose, target, steps=10):
    """Context number asked by the north star: the reference's own CUDA rasterizer (compiled unmodified,
    oracle/_ref) behind an eager-torch skinning chain on the same B200.  Not the reference arm."""
    try:
        import math
        import torch
        from oracle import ref_cuda
        from oracle import skin_oracle as so
        if not ref_cuda.available():
            return None
        dev = target.device
        if torch.cuda.current_stream(dev) != torch.cuda.default_stream(dev):
            return None
        T = lambda t: t.to(dev)
        leaf = lambda t: T(t).clone().requires_grad_(True)
        xyz, sc_raw, rot_raw, op_raw = leaf(scene.xyz), leaf(scene.scaling_raw), leaf(scene.rotation_raw), leaf(scene.opacity_raw)
        fdc, frest = leaf(scene.features_dc), leaf(scene.features_rest)
        joints, parents, mask = T(scene.joints), scene.parents, T(scene.motion_mask)
        nr = leaf(scene.node_radius_raw)
        ref = ref_cuda.RefRasterizer()
        W, H = cam.image_width, cam.image_height
        tfx, tfy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
        bg = torch.zeros(3, device=dev)
        view_m, proj_m, campos = cam.world_view_transform.contiguous(), cam.full_proj_transform.contiguous(), cam.camera_center.contiguous()

        def step():
            q = pose["local_rotation"].detach().clone().requires_grad_(True)
            g = pose["global_trans"].detach().clone().requires_grad_(True)
            d = so.deform_by_pose(xyz, joints, parents, nr, q, g, mask)
            means3D, scales, rots, opac = so.activate(xyz, sc_raw, rot_raw, op_raw, d["d_xyz"], d["d_rotation"])
            shs = torch.cat([fdc, frest], 1)
            a = [t.detach().contiguous() for t in (means3D, opac, shs, scales, rots)]
            f = ref.forward(bg, a[0], a[1], view_m, proj_m, campos, W, H, tfx, tfy, shs=a[2], scales=a[3], rotations=a[4])
            gimg = (torch.sign(f["color"] - target) / target.numel()).contiguous()
            z = torch.zeros(1, H, W, device=dev)
            gr = ref.backward(bg, a[0], f["radii"], f["alpha"], view_m, proj_m, campos, tfx, tfy, gimg, z, z, shs=a[2],
                              scales=a[3], rotations=a[4])
            torch.autograd.backward([means3D, scales, rots, opac, shs],
                                    [gr["dmean3D"], gr["dscale"], gr["drot"], gr["dopacity"], gr["dsh"]])
            for t in (xyz, sc_raw, rot_raw, op_raw, fdc, frest, nr):
                t.grad = None

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        t0 = time.time()
        for _ in range(steps):
            step()
        torch.cuda.synchronize()
        dt = (time.time() - t0) / steps
        return {"value": round(1.0 / dt, 3), "unit": UNIT, "ms_per_step": round(dt * 1e3, 3),
                "what": "eager-torch skinning on CUDA + the reference's CUDA rasterizer (oracle/_ref, unmodified sources) fwd+bwd"}
    except Exception as ex:  # context number only: never fail the bench for it
        return {"unavailable": repr(ex)[:200]}


def _emit(line: dict) -> None:
    """The ONE JSON line goes to the process's original stdout; see __main__."""
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


_JSON_FD = 1

if __name__ == "__main__":
    # Libraries write banners to file descriptor 1 behind Python's back (NCCL prints its version line there when the
    # first communicator is made): keep a private copy of stdout for the JSON line and point fd 1 at stderr.
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
