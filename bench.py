#!/usr/bin/env python
"""bench.py -- skin + splat-rasterize forward/backward iterations per second (BASELINE.json's metric).

    python bench.py --gpus N --steps K --warmup W            # our sm_100a path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port), host cores

A step = one pass of the hot path over one view: pose -> FK -> skin every Gaussian -> render -> L1 loss -> backward to
all Gaussian parameters and the skeleton inputs (train_rig.py:535-554 minus SSIM).  Default workload at every N: config
C0 of SURVEY.md section 8(d) = the configuration the metric is quoted on (300k Gaussians, 24-joint SMPL-like skeleton,
800x800), one view per GPU per step (weak scaling); camera, pose and target change every step (the scene's 8 views stay
resident in HBM and are visited round-robin); with N > 1 the Gaussian-parameter gradients are summed across the ranks
inside the timed region.

Other shapes of BASELINE.json's configs:
    --workload C3 --views-per-step 4      4-view batch per optimiser step (views sharded over the ranks, gradients
    --workload C4 --views-per-step 8      accumulated per rank, ONE all-reduce per step); iterations = views; strong scaling
    --workload C5 --mode sequence         256-frame novel-pose sequence, forward only, frames sharded over the ranks,
                                          no collective; metric = frames/s
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
N_VIEWS = 8  # views of the synthetic scene that the steps rotate over


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--mode", default="train", choices=["train", "sequence", "split"],
                    help="train: fwd+bwd steps (default); sequence: forward-only novel-pose frame sequence (C5)")
    ap.add_argument("--views-per-step", type=int, default=0,
                    help="global views per optimiser step, sharded over the ranks and accumulated per rank "
                         "(0 = one view per GPU per step, weak scaling)")
    ap.add_argument("--frame-sharding", default="interleaved", choices=["interleaved", "contiguous"],
                    help="sequence mode: frames r, r+N, ... per rank (balanced) or one contiguous range per rank")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--baselines", action="store_true",
                    help="also time BASELINE.md's B1 (CPU skinning alone, C1) and B2 (reference CUDA rasterizer alone) rows")
    ap.add_argument("--isotropic", action="store_true")
    ap.add_argument("--mlps", default="off", choices=["off", "on", "torch"],
                    help="on: the reference's demo configuration (scripts/run_demo.py:32): per-Gaussian skinning-weight MLP "
                         "and template-offset MLP evaluated (and differentiated) every step, on the tensor-core kernels; "
                         "torch: the same with the eager fp32 torch modules (what the reference runs)")
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
def algorithmic_bytes(P, R, HW, in_P=236, out_P=248, passes=6, n_grad=0, world=1, multimem=False):
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
    if world > 1:
        # bytes per NVLink direction and GPU: the switch-reduced kernel moves (N+1)/N of the buffer each way
        per["allreduce_multimem"] = int(n_grad * 4 * (world + 1) / world)
        # the peer kernel: a GPU receives (N-1)/N of the buffer through its own loads and (N-1)/N through its peers' stores
        per["allreduce_p2p"] = int(n_grad * 4 * 2 * (world - 1) / world)
    total = P * (2 * in_P + 336) + R * (148 + 24 * passes) + HW * 52
    own = sum(per[k] for k in ("skin_preprocess", "gauss_skin_backward", "depth_sort", "tile_count", "tile_scatter",
                               "blend_forward", "blend_backward")) + HW * 3 * 12
    return per, total, own


def make_scene(workload, isotropic=False, mode="train"):
    from riggs_b200 import synth
    if mode == "sequence":
        return synth.make_config(workload, isotropic=isotropic)
    return synth.make_config(workload, n_views=N_VIEWS, isotropic=isotropic)


def views_of_rank(rank, world, vps, n_views=N_VIEWS):
    """View indices this rank renders, step after step: step s covers global views s*world*vps .. (s+1)*world*vps - 1
    (mod n_views), rank r takes r, r + world, ... of them.  Returned: the flat list over one rotation period."""
    out, seen, s = [], set(), 0
    while True:
        grp = tuple((s * world * vps + rank + j * world) % n_views for j in range(vps))
        if grp in seen:
            break
        seen.add(grp)
        out.extend(grp)
        s += 1
    return out


def workload_string(name, scene, W, H, args, world):
    if args.mode == "split":
        return (f"{name}: {scene.N} Gaussians, {scene.J}-joint skeleton, {W}x{H}, ONE view per step rendered by all the GPUs "
                f"together (tile rows interleaved), fwd+bwd and forward only")
    if args.mode == "sequence":
        return (f"{name}: {scene.N} Gaussians, {scene.J}-joint skeleton, {W}x{H}, {len(scene.poses)}-frame novel-pose sequence, "
                f"forward only, frames sharded over the GPUs")
    if args.views_per_step > 0:
        return (f"{name}: {scene.N} Gaussians, {scene.J}-joint skeleton, {W}x{H}, {args.views_per_step} views/step sharded over "
                f"the GPUs, fwd+bwd, MLPs {args.mlps}")
    return f"{name}: {scene.N} Gaussians, {scene.J}-joint skeleton, {W}x{H}, 1 view/GPU/step, fwd+bwd, MLPs {args.mlps}"


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


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
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    if args.mode == "sequence":
        return run_sequence(args, rank, world, dev)
    if args.mode == "split":
        return run_split(args, rank, world, dev)

    V = args.views_per_step
    if V > 0 and V % world:
        raise SystemExit(f"--views-per-step {V} must be a multiple of the number of GPUs ({world})")
    vps = V // world if V > 0 else 1                      # views this rank renders per step
    views_global = V if V > 0 else world                  # iterations (views) per step over the whole job
    scene = make_scene(args.workload, args.isotropic)
    hp = HotPath(scene, device=dev, fused=not args.unfused, lambda_dssim=args.lambda_dssim,
                 mlps={"off": False, "on": True, "torch": "torch"}[args.mlps])
    order = views_of_rank(rank, world, vps)
    H, W = scene.cameras[0].image_height, scene.cameras[0].image_width
    g = torch.Generator().manual_seed(1000)
    targets_host_all = [torch.rand(3, H, W, generator=g).pin_memory() for _ in range(N_VIEWS)]
    cams = [scene.cameras[i].to(dev) for i in order]
    targets_host = [targets_host_all[i] for i in order]
    targets = [t.to(dev) for t in targets_host]
    poses_host = [{k: v.clone().pin_memory() for k, v in scene.poses[i].items()} for i in order]
    poses = [{k: v.to(dev) for k, v in p.items()} for p in poses_host]
    n_groups = len(order) // vps
    params = hp.parameters()
    n_grad = sum(p.numel() for p in params)
    if world > 1 or vps > 1:
        # the backward kernels write (first view) / add (later views) into one flat fp32 buffer, summed over the ranks in place
        hp.enable_flat_grads()

    def reduce_grads():
        if world > 1:
            hp.reduce_grads()

    def group(i):
        j = (i % n_groups) * vps
        return cams[j:j + vps], poses[j:j + vps], targets[j:j + vps]

    def eager_step(i):
        hp.zero_grad()
        cs, ps, ts = group(i)
        ps = [{k: v.detach().clone().requires_grad_(True) for k, v in p.items()} for p in ps]
        loss = hp.step_views(cs, ps, ts) if vps > 1 else hp.step(cs[0], ps[0], ts[0])
        reduce_grads()
        return loss

    # ---- the all-reduce kernel's result against torch.distributed's, once, outside the timed regions -----------
    allreduce_checked = None
    if world > 1:
        eager_step(0)  # leaves every rank's own gradients reduced; now redo it with a reference copy
        hp.zero_grad()
        cs, ps, ts = group(0)
        ps = [{k: v.detach().clone().requires_grad_(True) for k, v in p.items()} for p in ps]
        overlap_mode, hp.flat.overlap = hp.flat.overlap, False  # this step must leave the rank's OWN gradients in the buffer
        (hp.step_views(cs, ps, ts) if vps > 1 else hp.step(cs[0], ps[0], ts[0]))
        ref = hp.flat.flat.clone()
        dist.all_reduce(ref, op=dist.ReduceOp.SUM)
        hp.reduce_grads()
        torch.cuda.synchronize()
        hp.flat.overlap = overlap_mode
        if getattr(hp.flat, "_mc_ptr", 0):  # switch-side sum: the order of the adds is the switch's
            err = float((hp.flat.flat - ref).abs().max() / ref.abs().max().clamp_min(1e-30))
            ok = err <= 1e-6
        elif getattr(hp.flat, "hdl", None) is not None:  # rank-ordered adds on every rank: bit-identical to each other
            err = float((hp.flat.flat - ref).abs().max() / ref.abs().max().clamp_min(1e-30))
            ok = err <= 1e-6
            chk = hp.flat.flat.clone()
            dist.broadcast(chk, src=0)
            ok = ok and bool(torch.equal(chk, hp.flat.flat))
        else:
            err, ok = 0.0, True
        err_overlap = None
        if overlap_mode:  # the sum taken chunk by chunk while the backward runs, on the same inputs (float atomics: not bit-equal)
            eager_step(0)
            torch.cuda.synchronize()
            err_overlap = float((hp.flat.flat - ref).abs().max() / ref.abs().max().clamp_min(1e-30))
            ok = ok and err_overlap <= 1e-4
        t_ok = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        allreduce_checked = {"ok": bool(int(t_ok[0])), "max_rel_err_vs_nccl": err}
        if err_overlap is not None:
            allreduce_checked["overlapped_with_backward_max_rel_err"] = err_overlap
        if not allreduce_checked["ok"]:
            raise SystemExit(f"rank {rank}: own all-reduce disagrees with dist.all_reduce (rel err {err})")

    # The fused step is captured as CUDA graphs (riggs_b200.hotpath.RotatingSteps: one graph per group of views, the
    # views' inputs resident in HBM) and replayed round-robin; the cross-rank sum stays an eager call after each replay.
    use_graph = hp.fused and not args.no_graph
    rot = None
    if use_graph:
        from riggs_b200.hotpath import RotatingSteps
        rot = RotatingSteps(hp, cams, poses, targets, views_per_step=vps, warmup=2)

    step_no = [0]

    def one_step():
        i = step_no[0]
        step_no[0] += 1
        if rot is None:
            return eager_step(i)
        loss = rot.step()  # the next group's graph: camera / pose / target change, nothing is copied
        reduce_grads()
        return loss

    L = _lib.lib()
    # nvidia-smi takes a driver lock while it starts: start it before the warm-up, not inside the timed region
    clocks = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        one_step()
    torch.cuda.synchronize()
    if clocks is not None:
        t_wait = time.time()
        while not clocks.rows and time.time() - t_wait < 5.0:
            time.sleep(0.05)
    for _ in range(3):  # every rank runs the same number of steps (each one contains a collective when N > 1)
        one_step()
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
        one_step()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.time()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    overflowed = bool(rot.overflowed) if rot is not None else False
    # kernels of this library inside the timed region (a replayed graph launches what its capture recorded)
    launches = rot.launches_per_replay * args.steps if rot is not None else L.rg_launch_count() - launches0
    if world > 1 and rot is not None and not rot.slots[0]._reduces_in_graph:
        launches += args.steps  # the all-reduce kernel, launched eagerly after each replay (NCCL fallback: not ours, not counted)
    clk = clocks.stop(t0, t1) if clocks else None

    # ---- per-kernel durations: a separate eager pass with every launch bracketed by CUDA events on its stream ----
    L.rg_profile_reset()
    L.rg_profile_enable(1)
    for i in range(args.steps):
        eager_step(i)
    torch.cuda.synchronize()
    L.rg_profile_enable(0)
    prof = _lib.profile_read()
    L.rg_profile_reset()
    if world > 1:
        dist.barrier()

    # ---- timed region 2: end to end through the public API, host buffers ------------------------------
    class _HostCam:  # pinned host copies of one view's camera tensors
        def __init__(self, c):
            self.world_view_transform = c.world_view_transform.cpu().pin_memory()
            self.full_proj_transform = c.full_proj_transform.cpu().pin_memory()
            self.camera_center = c.camera_center.cpu().pin_memory()

    cams_host = [_HostCam(c) for c in cams]

    def host_group(i):
        j = (i % n_groups) * vps
        c, p, t = cams_host[j:j + vps], poses_host[j:j + vps], targets_host[j:j + vps]
        return (c, p, t) if vps > 1 else (c[0], p[0], t[0])

    h2d = vps * (targets_host[0].numel() * 4 + sum(v.numel() * 4 for v in poses_host[0].values()) + (16 + 16 + 3) * 4)
    pipe = None
    if use_graph:  # two captured graphs alternate so that step i+1's inputs land while step i runs (set-up, untimed)
        from riggs_b200.hotpath import PipelinedSteps
        c0, p0, t0_ = group(0)
        first = (c0, p0, t0_) if vps > 1 else (c0[0], p0[0], t0_[0])
        pipe = PipelinedSteps(hp, *first, warmup=2, pool=rot.pool)
        for i in range(2):
            pipe.feed(*host_group(i))
            pipe.result(pipe.launch())
            reduce_grads()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    f0.record()
    last = 0.0
    if pipe is not None:  # input pipeline: step i+1's host -> device copies run while step i computes
        pipe.feed(*host_group(0))
    for it in range(args.steps):
        if pipe is not None:
            pipe.launch()  # waits for this step's inputs, replays, queues the read-back of the step's loss
            if it + 1 < args.steps:
                pipe.feed(*host_group(it + 1))
            reduce_grads()
            if it > 0:
                last = pipe.result(it - 1)  # host read of the previous step's loss while this step runs
            continue
        c, p, t = host_group(it)
        c, p, t = (c, p, t) if vps > 1 else ([c], [p], [t])
        import copy
        dc = []
        for hc, src in zip(c, group(it)[0]):
            cc = copy.copy(src)
            cc.world_view_transform = hc.world_view_transform.to(dev, non_blocking=True)
            cc.full_proj_transform = hc.full_proj_transform.to(dev, non_blocking=True)
            cc.camera_center = hc.camera_center.to(dev, non_blocking=True)
            dc.append(cc)
        dp = [{k: v.to(dev, non_blocking=True).requires_grad_(True) for k, v in pp.items()} for pp in p]
        dt = [tt.to(dev, non_blocking=True) for tt in t]
        hp.zero_grad()
        loss = hp.step_views(dc, dp, dt) if vps > 1 else hp.step(dc[0], dp[0], dt[0])
        reduce_grads()
        last = loss.detach().cpu()  # device -> host read of the step's result
    if pipe is not None:
        last = pipe.result(args.steps - 1)
    f1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_e2e = f0.elapsed_time(f1)
    last = float(last) if not hasattr(last, "numel") else float(last.reshape(-1)[-1])
    overflowed = overflowed or bool(pipe is not None and pipe.overflowed_steps)

    # NVLink: a peer copy of the gradient buffer's size on the same box, for the all-reduce's achieved-vs-link figure
    nvlink = peer_copy_bandwidth(hp, dev) if world > 1 else None

    # max over ranks
    per_rank_ms = None
    if world > 1:
        mine = torch.tensor([ms / args.steps], device=dev)
        allms = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allms, mine)
        per_rank_ms = [round(float(x), 4) for x in allms]
        t = torch.tensor([ms, ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # workload facts for the roofline (first view of this rank)
    with torch.no_grad():
        out = hp.forward(cams[0], poses[0])
    torch.cuda.synchronize()
    P = scene.N
    vis = int((out["radii"] > 0).sum())
    R = int(out["status"][0]) if out.get("status") is not None else _count_rendered(hp, cams[0], poses[0])
    in_P = 236 - (8 if args.isotropic else 0)
    multimem = bool(getattr(hp.flat, "_mc_ptr", 0)) if hp.flat is not None else False
    per, total_bytes, own_bytes = algorithmic_bytes(P, R, H * W, in_P=in_P, n_grad=n_grad, world=world, multimem=multimem)
    step_ms = ms / args.steps
    view_ms = step_ms / vps
    value = views_global * args.steps / (ms / 1e3)
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured" if "hbm_gbs" in peaks else "fallback"
    kern = {k: {"ms_per_launch": v[0] / max(v[1], 1), "launches": v[1], "ms_per_step": v[0] / args.steps}
            for k, v in prof.items()}
    dom = max(kern, key=lambda k: kern[k]["ms_per_step"]) if kern else None
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the `ncu --set full` capture of this workload
    # (profiles/r2_*_full_raw.csv, C0 on one B200); other workloads have no capture -> null
    ncu_traffic = NCU_TRAFFIC if args.workload == "C0" and not args.isotropic and args.mlps == "off" else {}
    if args.mlps != "off":  # the MLP kernels are tensor-core work: FLOPs, not bytes
        J = scene.J
        macs = lambda e, o: e * 256 + 4 * 256 * 256 + (e + 256) * 256 + 2 * 256 * 256 + 256 * o
        fl_w, fl_d = 2 * P * macs(63, J - 1), 2 * P * macs(27, 3)
        per["mlp_forward"] = (fl_w + fl_d) // 2  # per launch: one of the two MLPs (their mean)
        per["mlp_backward_chain"] = (fl_w + fl_d) // 2
        per["mlp_wgrad"] = (fl_w + fl_d) // 2
    roof = None
    if dom and dom.startswith("mlp_"):
        fl = per.get(dom, 0)
        tf = fl / (kern[dom]["ms_per_launch"] * 1e-3) / 1e12
        peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
        roof = {"kernel": dom, "bound": "tensor", "achieved": round(tf, 2), "peak": peak_tf, "unit": "TFLOP/s",
                "frac": round(tf / peak_tf, 4), "traffic": None,
                "peak_source": "measured (sustained bf16 GEMM)" if "bf16_tflops_sustained" in peaks else "fallback",
                "algorithmic_flops_per_launch": fl, "ms_per_launch": round(kern[dom]["ms_per_launch"], 4)}
    elif dom:
        b = per.get(dom, 0)
        achieved = b / (kern[dom]["ms_per_launch"] * 1e-3) / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": ncu_traffic.get(dom), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": b, "ms_per_launch": round(kern[dom]["ms_per_launch"], 4)}
        if dom.startswith("allreduce") and nvlink:
            roof.update(bound="nvlink", peak=nvlink["peer_read_gbs"], frac=round(achieved / max(nvlink["peer_read_gbs"], 1e-9), 4),
                        peak_source="peer copy of the same size measured in this run")
    footprint = P * (in_P + 248) + own_bytes - per["skin_preprocess"] - per["gauss_skin_backward"] + P * 150
    par = "single GPU"
    if world > 1:
        kind = ("own all-reduce kernel over symmetric memory (" + ("rg_allreduce_multimem, NVLS" if multimem else
                "rg_allreduce_p2p, NVLink peer loads/stores") + ")") if getattr(hp.flat, "hdl", None) is not None else "NCCL all-reduce"
        par = f"dp{world} over views, {kind} of {n_grad} grad floats"
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(step_ms, 4), "higher_is_better": True,
        "scaling": "strong" if V > 0 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, scene, W, H, args, world), "P": P, "visible": vis, "R": R,
                   "HW": H * W, "views_per_step_per_gpu": vps,
                   "views": f"{len(order)} resident views visited round-robin (camera, pose and target change every step)",
                   "l2": (f"no explicit flush: a step reads the parameters and writes their gradients once "
                          f"(~{footprint >> 20} MB touched per view incl. the per-frame state) against a 126 MB L2, and the "
                          f"view changes every step; the per-frame state itself (~{(R * 5 + P * 150 + H * W * 28) >> 20} MB) "
                          f"stays L2-resident between the kernels of one step (ncu: dram bytes << algorithmic bytes for the blend kernels)"),
                   "launch": "cuda graph replay" if rot is not None else "eager",
                   "loss": "L1" if args.lambda_dssim <= 0 else f"(1-{args.lambda_dssim})*L1 + {args.lambda_dssim}*(1-SSIM)",
                   "mlps": args.mlps, "parallelism": par},
        "e2e": {"value": round(views_global * args.steps / (ms_e2e / 1e3), 3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4 * vps, "last_loss": last},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": roof,
        "step_hbm": {"algorithmic_bytes_per_view": total_bytes, "achieved_gbs": round(total_bytes / (view_ms * 1e-3) / 1e9, 2),
                     "frac_of_peak": round(total_bytes / (view_ms * 1e-3) / 1e9 / peak, 4), "peak": peak,
                     "own_bytes_per_view": own_bytes,
                     "own_frac_of_peak": round(own_bytes / (view_ms * 1e-3) / 1e9 / peak, 4),
                     "note": "algorithmic = SURVEY 8(d) contract formula (charges a 6-pass 64-bit key sort this design does not run); "
                             "own = the bytes this design's kernels are asked to move"},
        "kernels": {k: {kk: round(vv, 4) if isinstance(vv, float) else vv for kk, vv in v.items()} for k, v in kern.items()},
        "instance_buffer_overflowed": overflowed,
    }
    if allreduce_checked is not None:
        line["allreduce_checked"] = allreduce_checked
    if per_rank_ms is not None:
        line["per_rank_ms_per_step"] = per_rank_ms
    if nvlink:
        ar = [k for k in kern if k.startswith("allreduce")]
        if ar:
            k = ar[0]
            nvlink["allreduce_kernel"] = k
            nvlink["allreduce_ms"] = round(kern[k]["ms_per_launch"], 4)
            nvlink["allreduce_gbs_per_direction"] = round(per.get(k, 0) / (kern[k]["ms_per_launch"] * 1e-3) / 1e9, 1)
        line["nvlink"] = nvlink
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(scene, args)
    if world == 1 and not args.no_ref_cuda:
        rc = ref_cuda_baseline(scene, cams[0], poses[0], targets[0], steps=min(args.steps, 10), mlps=hp if args.mlps != "off" else None)
        if rc:
            line["reference_cuda_b200"] = rc
    if world == 1 and args.baselines:
        line["reference_cpu_skinning"] = reference_cpu_skinning()
        line["reference_cuda_rasterizer"] = reference_cuda_rasterizer(["C2", "C3", "C4", "C5"])
    if world == 1:
        line["next_rows"] = next_rows(hp, out["render"].detach(), targets[0])
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, C0 on one B200 (profiles/r2_r_full_raw.csv)
NCU_TRAFFIC = {"blend_backward": 17107968, "blend_forward": 14824192, "gauss_skin_backward": 130815488,
               "skin_preprocess": 79057408, "tile_scatter": 18270976, "tile_count": 6019840}


def peer_copy_bandwidth(hp, dev, iters=10):
    """GB/s of a plain device copy that READS a peer's copy of the gradient buffer over NVLink (what one direction of
    one link pair delivers to a copy of this size on this box)."""
    import torch
    flat = hp.flat
    hdl = getattr(flat, "hdl", None)
    if hdl is None:
        return None
    try:
        n = flat._storage.numel()
        peer = (hdl.rank + 1) % hdl.world_size
        src = hdl.get_buffer(peer, (n,), torch.float32)
        dst = torch.empty(n, dtype=torch.float32, device=dev)
        hdl.barrier(channel=0)
        for _ in range(2):
            dst.copy_(src)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            dst.copy_(src)
        e1.record()
        torch.cuda.synchronize()
        hdl.barrier(channel=0)
        msc = e0.elapsed_time(e1) / iters
        return {"peer_read_gbs": round(n * 4 / (msc * 1e-3) / 1e9, 1), "bytes": n * 4,
                "how": "torch copy_ from the next rank's symmetric-memory buffer, CUDA events, all ranks at once"}
    except Exception as ex:
        return {"unavailable": repr(ex)[:160]}


# ---------------------------------------------------------------------------------------------------------
def run_sequence(args, rank, world, dev):
    """C5 as BASELINE.json states it: a 256-frame novel-pose sequence, forward only, frames sharded contiguously over
    the ranks, no collective in the timed region (model: render_rig.py:250-334)."""
    import torch
    import torch.distributed as dist
    from riggs_b200 import _lib
    from riggs_b200.dist import shard_frames, shard_frames_interleaved
    from riggs_b200.hotpath import HotPath, PoseSequenceRenderer

    scene = make_scene(args.workload, args.isotropic, mode="sequence")
    hp = HotPath(scene, device=dev, fused=True, mlps={"off": False, "on": True, "torch": "torch"}[args.mlps], inference=True)
    cam = scene.cameras[0].to(dev)
    H, W = cam.image_height, cam.image_width
    F_total = len(scene.poses)
    if args.frame_sharding == "contiguous":
        b, e = shard_frames(F_total, rank, world)
        mine = list(range(b, e))
    else:
        mine = shard_frames_interleaved(F_total, rank, world)
    F = len(mine)
    rot_host = torch.stack([scene.poses[f]["local_rotation"] for f in mine]).pin_memory()
    trans_host = torch.stack([scene.poses[f]["global_trans"] for f in mine]).pin_memory()
    seq = PoseSequenceRenderer(hp, cam, scene.J, max_frames=max(F, 1))
    seq.load(rot_host, trans_host)
    torch.cuda.synchronize()
    L = _lib.lib()

    def one_pass():
        seq.rewind()
        for _ in range(F):
            seq.render_next()

    clocks = ClockSampler(dev.index or 0) if rank == 0 else None
    for _ in range(max(min(args.warmup, 3), 1)):
        one_pass()
    torch.cuda.synchronize()
    if clocks is not None:
        t_wait = time.time()
        while not clocks.rows and time.time() - t_wait < 5.0:
            time.sleep(0.05)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        one_pass()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.time()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop(t0, t1) if clocks else None
    overflowed = seq.overflowed

    # per-kernel durations of one eager frame loop
    L.rg_profile_reset()
    L.rg_profile_enable(1)
    n_prof = min(F, 16)
    with torch.no_grad():
        for f in range(n_prof):
            hp.forward(cam, {"local_rotation": seq.rot[f], "global_trans": seq.trans[f]})
    torch.cuda.synchronize()
    L.rg_profile_enable(0)
    prof = _lib.profile_read()
    L.rg_profile_reset()

    # end to end: the pose sequence comes from pinned host memory every pass, every frame's image goes to pinned host memory
    ring = [torch.empty(3, H, W).pin_memory() for _ in range(2)]
    out_stream = torch.cuda.Stream(device=dev)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    f0.record()
    for _ in range(args.steps):
        seq.load(rot_host, trans_host)
        for _ in range(F):
            slot, img = seq.render_next()
            done = torch.cuda.Event()
            done.record()
            with torch.cuda.stream(out_stream):
                out_stream.wait_event(done)
                ring[slot].copy_(img, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(out_stream)
            seq.release(slot, ev)
    torch.cuda.current_stream().wait_stream(out_stream)
    f1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms_e2e = f0.elapsed_time(f1)
    per_rank_ms = None
    if world > 1:
        mine = torch.tensor([ms / (args.steps * max(F, 1))], device=dev)
        allms = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allms, mine)
        per_rank_ms = [round(float(x), 4) for x in allms]
        t = torch.tensor([ms, ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    with torch.no_grad():
        out = hp.forward(cam, {"local_rotation": seq.rot[0], "global_trans": seq.trans[0]})
    torch.cuda.synchronize()
    P, R = scene.N, int(out["status"][0])
    vis = int((out["radii"] > 0).sum())
    in_P = 236 - (8 if args.isotropic else 0)
    fwd_bytes = P * (in_P + 48) + R * 208 + H * W * 24       # B_fwd
    own_fwd = P * (in_P + 60) + P * 64 + P * 40 + R * 4 + R * 44 + H * W * 24
    frames = F_total * args.steps
    frame_ms = ms / (args.steps * max(F, 1))
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    kern = {k: {"ms_per_launch": v[0] / max(v[1], 1), "launches": v[1], "ms_per_frame": v[0] / n_prof} for k, v in prof.items()}
    dom = max(kern, key=lambda k: kern[k]["ms_per_frame"]) if kern else None
    per = {"skin_preprocess": P * (in_P + 60), "blend_forward": R * 44 + H * W * 24, "depth_sort": P * 64,
           "tile_count": P * 20, "tile_scatter": P * 20 + R * 4, "tile_offsets": 0}
    roof = None
    if dom:
        bts = per.get(dom, 0)
        ach = bts / (kern[dom]["ms_per_launch"] * 1e-3) / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": round(ach, 2), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                "traffic": None, "algorithmic_bytes_per_launch": bts, "ms_per_launch": round(kern[dom]["ms_per_launch"], 4),
                "peak_source": "measured" if "hbm_gbs" in peaks else "fallback"}
    line = {
        "metric": "novel-pose frames/s (skin + rasterize, forward only)", "value": round(frames / (ms / 1e3), 3), "unit": "frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(min(args.warmup, 3), 1), "ms_per_step": round(ms / args.steps, 4),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, scene, W, H, args, world), "P": P, "visible": vis, "R_frame0": R,
                   "HW": H * W, "frames_total": F_total, "frames_per_gpu": F, "ms_per_frame_per_gpu": round(frame_ms, 4),
                   "step": "one pass over the whole sequence", "mlps": args.mlps,
                   "l2": f"no explicit flush: one frame reads {P * in_P >> 20} MB of parameters (> 126 MB L2) and the pose changes every frame",
                   "launch": "cuda graph replay, the graph advances its own frame index", "parallelism":
                       (f"frames sharded over {world} GPUs ({args.frame_sharding}), no collective" if world > 1 else "single GPU")},
        "e2e": {"value": round(frames / (ms_e2e / 1e3), 3), "unit": "frames/s",
                "h2d_bytes_per_step": int(rot_host.numel() * 4 + trans_host.numel() * 4), "d2h_bytes_per_step": int(F * 3 * H * W * 4)},
        "gpu_launches": int(seq.launches_per_replay * F * args.steps),
        "clocks": clk, "roofline": roof,
        "step_hbm": {"algorithmic_bytes_per_frame": fwd_bytes, "frac_of_peak": round(fwd_bytes / (frame_ms * 1e-3) / 1e9 / peak, 4),
                     "own_bytes_per_frame": own_fwd, "own_frac_of_peak": round(own_fwd / (frame_ms * 1e-3) / 1e9 / peak, 4), "peak": peak},
        "kernels": {k: {kk: round(vv, 4) if isinstance(vv, float) else vv for kk, vv in v.items()} for k, v in kern.items()},
        "instance_buffer_overflowed": bool(overflowed),
    }
    if per_rank_ms is not None:
        line["per_rank_ms_per_frame"] = per_rank_ms
    _emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_split(args, rank, world, dev):
    """Latency of ONE view when all the GPUs work on it (dist.ViewSplit): the training step (fwd + L1 + bwd; every rank
    ends with the full gradients, no gradient all-reduce) and the forward-only frame.  `value` = steps/s of the whole job
    (one view per step, whatever N is: strong scaling)."""
    import torch
    import torch.distributed as dist
    from riggs_b200 import _lib
    from riggs_b200.hotpath import HotPath, PoseSequenceRenderer, RotatingSteps

    scene = make_scene(args.workload, args.isotropic)
    hp = HotPath(scene, device=dev, fused=True)
    H, W = scene.cameras[0].image_height, scene.cameras[0].image_width
    order = list(range(N_VIEWS))
    g = torch.Generator().manual_seed(1000)
    targets_host = [torch.rand(3, H, W, generator=g).pin_memory() for _ in range(N_VIEWS)]
    cams = [scene.cameras[i].to(dev) for i in order]
    targets = [t.to(dev) for t in targets_host]
    poses_host = [{k: v.clone().pin_memory() for k, v in scene.poses[i].items()} for i in order]
    poses = [{k: v.to(dev) for k, v in p.items()} for p in poses_host]
    params = hp.parameters()

    def eager(i, split):
        old, hp.split = hp.split, split
        try:
            hp.zero_grad()
            pose = {k: v.detach().clone().requires_grad_(True) for k, v in poses[i].items()}
            out = hp.forward(cams[i], pose)
            from riggs_b200.loss_utils import l1_loss_and_grad
            loss, grad = l1_loss_and_grad(out["render"], targets[i])
            out["render"].backward(grad)
            torch.cuda.synchronize()
            # nothing that carries a grad_fn leaves this function: an autograd graph kept alive breaks the later captures
            res = {k: out[k].detach().clone() for k in ("render", "depth", "alpha", "radii")}
            return res, [p.grad.clone() for p in params] + [pose["local_rotation"].grad.clone()], int(out["status"][0])
        finally:
            hp.split = old

    # ---- the split view against the same view rendered by this rank alone, once, outside the timed regions ----
    split = hp.enable_view_split(W, H)
    checked = None
    if split is not None:
        hp.split = None
        whole, g_whole, n_whole = eager(0, None)
        _, g_whole2, _ = eager(0, None)
        part, g_part, n_part = eager(0, split)
        rel = lambda x, y: max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)) for a, b in zip(x, y))
        spread = rel(g_whole2, g_whole)  # the unsplit step against itself: float atomics land in another order every run
        same = all(bool(torch.equal(part[k], whole[k])) for k in ("render", "depth", "alpha", "radii"))
        err = rel(g_part, g_whole)
        chk = g_part[0].clone()
        dist.broadcast(chk, src=0)
        identical = bool(torch.equal(chk, g_part[0]))  # replicated backward on bit-identical sums: same bits on every rank
        cnt = torch.tensor([n_part], device=dev)
        dist.all_reduce(cnt)
        ok = same and err <= min(max(5e-5, 4.0 * spread), 1e-3) and int(cnt[0]) == n_whole
        t_ok = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
        checked = {"ok": bool(int(t_ok[0])), "image_bit_identical_to_unsplit": same, "max_grad_rel_err_vs_unsplit": err,
                   "unsplit_run_to_run_rel_err": spread, "gradients_identical_on_all_ranks": identical, "instances_unsplit": n_whole, "instances_this_rank": n_part,
                   "peer_memory": split.images.hdl is not None}
        if not checked["ok"]:
            raise SystemExit(f"rank {rank}: split view disagrees with the unsplit view: {checked}")
        hp.split = split
        del whole, part, g_whole, g_part, g_whole2
    hp.zero_grad()

    rot = RotatingSteps(hp, cams, poses, targets, views_per_step=1, warmup=2)
    L = _lib.lib()
    clocks = ClockSampler(dev.index or 0) if rank == 0 else None
    for _ in range(max(args.warmup, 3) + 3):
        rot.step()
    torch.cuda.synchronize()
    if clocks is not None:
        t_wait = time.time()
        while not clocks.rows and time.time() - t_wait < 5.0:
            time.sleep(0.05)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        rot.step()
    e1.record()
    torch.cuda.synchronize()
    t1 = time.time()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop(t0, t1) if clocks else None
    overflowed = bool(rot.overflowed)

    # per-kernel durations (eager pass)
    L.rg_profile_reset()
    L.rg_profile_enable(1)
    for i in range(args.steps):
        eager(i % N_VIEWS, hp.split)
    torch.cuda.synchronize()
    L.rg_profile_enable(0)
    prof = _lib.profile_read()
    L.rg_profile_reset()

    # forward only: the same views as a frame sequence (graph replay, all ranks on every frame)
    seq = PoseSequenceRenderer(hp, cams[0], scene.J, max_frames=N_VIEWS)
    seq.load(torch.stack([p["local_rotation"] for p in poses]), torch.stack([p["global_trans"] for p in poses]))
    n_fr = max(args.steps, N_VIEWS) // N_VIEWS * N_VIEWS
    for _ in range(N_VIEWS):
        seq.render_next()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for f in range(n_fr):
        if f % N_VIEWS == 0:
            seq.rewind()
        seq.render_next()
    s1.record()
    torch.cuda.synchronize()
    ms_fwd = s0.elapsed_time(s1) / n_fr

    # end to end: pose and target from pinned host memory every step, the loss read back
    from riggs_b200.hotpath import PipelinedSteps
    pipe = PipelinedSteps(hp, cams[0], poses[0], targets[0], warmup=2, pool=rot.pool)
    for i in range(2):
        pipe.feed(None, poses_host[i], targets_host[i])
        pipe.result(pipe.launch())
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    pipe.feed(None, poses_host[0], targets_host[0])
    last = 0.0
    for it in range(args.steps):
        pipe.launch()
        if it + 1 < args.steps:
            pipe.feed(None, poses_host[(it + 1) % N_VIEWS], targets_host[(it + 1) % N_VIEWS])
        if it > 0:
            last = pipe.result(it - 1)
    last = pipe.result(args.steps - 1)
    f1.record()
    torch.cuda.synchronize()
    ms_e2e = f0.elapsed_time(f1)
    per_rank = None
    if world > 1:
        mine = torch.tensor([ms / args.steps], device=dev)
        allms = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allms, mine)
        per_rank = [round(float(x), 4) for x in allms]
        t = torch.tensor([ms, ms_e2e, ms_fwd], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e, ms_fwd = float(t[0]), float(t[1]), float(t[2])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    P = scene.N
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    kern = {k: {"ms_per_launch": v[0] / max(v[1], 1), "launches": v[1], "ms_per_step": v[0] / args.steps} for k, v in prof.items()}
    dom = max(kern, key=lambda k: kern[k]["ms_per_step"]) if kern else None
    in_P = 236 - (8 if args.isotropic else 0)
    R = checked["instances_unsplit"] if checked else int(rot.slots[0]._ws.status[0])
    per, total_bytes, own_bytes = algorithmic_bytes(P, R, H * W, in_P=in_P)
    roof = None
    if dom:
        # a rank's blend / binning kernels see 1/N of the instances and pixels; the per-Gaussian kernels all of P
        shard = world if dom in ("blend_backward", "blend_forward", "tile_scatter") else 1
        b = per.get(dom, 0) // shard
        if dom.startswith("allreduce"):
            b = (10 * P + 5 * H * W) * 4 // 2  # mean of the two sums' payloads per launch
        ach = b / (kern[dom]["ms_per_launch"] * 1e-3) / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": round(ach, 2), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                "traffic": None, "algorithmic_bytes_per_launch": b, "ms_per_launch": round(kern[dom]["ms_per_launch"], 4),
                "peak_source": "measured" if "hbm_gbs" in peaks else "fallback"}
    step_ms = ms / args.steps
    line = {
        "metric": METRIC, "value": round(args.steps / (ms / 1e3), 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(step_ms, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(args.workload, scene, W, H, args, world), "P": P, "R": R, "HW": H * W,
                   "views": f"{N_VIEWS} resident views visited round-robin", "launch": "cuda graph replay",
                   "l2": "no explicit flush: the view changes every step and a step reads the parameters and writes their gradients once",
                   "parallelism": (f"one view over {world} GPUs: tile row y on GPU y % {world}; skin / project / depth sort and the "
                                   f"per-Gaussian backward replicated; two in-graph sums per step (image rows {5 * H * W * 4 >> 10} KB, "
                                   f"per-Gaussian screen-space accumulators {10 * P * 4 >> 10} KB); no gradient all-reduce")
                   if world > 1 else "single GPU"},
        "forward_only_ms_per_frame": round(ms_fwd, 4),
        "e2e": {"value": round(args.steps / (ms_e2e / 1e3), 3), "unit": UNIT,
                "h2d_bytes_per_step": int(targets_host[0].numel() * 4 + sum(v.numel() * 4 for v in poses_host[0].values())),
                "d2h_bytes_per_step": 4, "last_loss": float(last)},
        "gpu_launches": int(rot.launches_per_replay * args.steps), "clocks": clk, "roofline": roof,
        "kernels": {k: {kk: round(vv, 4) if isinstance(vv, float) else vv for kk, vv in v.items()} for k, v in kern.items()},
        "instance_buffer_overflowed": overflowed,
    }
    if checked is not None:
        line["split_checked"] = checked
    if per_rank is not None:
        line["per_rank_ms_per_step"] = per_rank
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
_CPU_MLPS = {}


def _cpu_mlps(scene):
    """The two per-Gaussian MLPs on the host (modules pinned to the reference's own, tests/test_mlp_front_end.py),
    with the seeded weights the GPU arm uses (HotPath(mlps=True))."""
    import torch
    from riggs_b200 import skeleton_warp as sw
    key = (scene.N, scene.J)
    if key not in _CPU_MLPS:
        torch.manual_seed(4321)
        wm = sw.WeightMLP(3, scene.J - 1)
        dn = sw.DeformMLP(4 * scene.J)
        with torch.no_grad():
            dn.gaussian_warp.weight.normal_(0, 1e-3)
        _CPU_MLPS[key] = (wm, dn)
    return _CPU_MLPS[key]


def cpu_step(scene, view, target, mlps=False):
    """One full iteration on the host cores; returns the loss."""
    import numpy as np
    import torch
    from oracle import oracle_py as orc
    from oracle import skin_oracle as so
    import math
    cam = scene.cameras[view % len(scene.cameras)]
    p = scene.poses[view % len(scene.poses)]
    leaf = lambda t: t.clone().requires_grad_(True)
    xyz, sc_raw, rot_raw, op_raw = leaf(scene.xyz), leaf(scene.scaling_raw), leaf(scene.rotation_raw), leaf(scene.opacity_raw)
    fdc, frest = leaf(scene.features_dc), leaf(scene.features_rest)
    q, g, nr = leaf(p["local_rotation"]), leaf(p["global_trans"]), leaf(scene.node_radius_raw)
    mlp_w = offs = None
    if mlps:
        wm, dn = _cpu_mlps(scene)
        for m in (wm, dn):
            for prm in m.parameters():
                prm.grad = None
        mlp_w = wm(xyz.detach())
        offs = dn(xyz.detach(), q.detach().reshape(-1)[None].expand(scene.N, -1))
    d = so.deform_by_pose(xyz, scene.joints, scene.parents, nr, q, g, scene.motion_mask, mlp_w, offs)
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


def _cpu_targets(scene):
    import torch
    cam = scene.cameras[0]
    g = torch.Generator().manual_seed(1000)
    return [torch.rand(3, cam.image_height, cam.image_width, generator=g) for _ in range(N_VIEWS)]


def cpu_baseline(scene, args, budget_s=12.0):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    targets = _cpu_targets(scene)
    mlps = args.mlps != "off"
    cpu_step(scene, 0, targets[0], mlps)  # warm-up (also builds the oracle library)
    n, t0 = 0, time.time()
    while True:
        cpu_step(scene, n, targets[n % N_VIEWS], mlps)
        n += 1
        if time.time() - t0 > budget_s or n >= 40:  # a bounded sample: ~10-15 s of host work
            break
    dt = (time.time() - t0) / n
    return {"value": round(1.0 / dt, 4), "unit": UNIT, "cores": cores, "kind": "port", "cpu": cpu_model(),
            "sample": f"{n} full {args.workload} iterations (skin_oracle.py torch-CPU + raster_oracle.c OpenMP"
                      f"{', both MLPs in torch-CPU' if mlps else ''}), {dt:.2f} s each"}


def run_reference(args):
    """The reference's CPU implementation of the path, as its oracle port, on all host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    scene = make_scene(args.workload, args.isotropic, mode=args.mode)
    cam = scene.cameras[0]
    W, H = cam.image_width, cam.image_height
    if args.mode == "sequence":
        return run_reference_sequence(args, scene, world, cores)
    targets = _cpu_targets(scene)
    mlps = args.mlps != "off"
    warm = max(args.warmup, 3)
    t_w = time.time()
    for i in range(warm):
        cpu_step(scene, i, targets[i % N_VIEWS], mlps)
        if time.time() - t_w > 60:  # the warm-up is part of the contract, but a slow host must not eat the run
            warm = i + 1
            break
    t0 = time.time()
    steps = 0
    for i in range(args.steps):
        cpu_step(scene, i, targets[i % N_VIEWS], mlps)
        steps += 1
        if time.time() - t0 > 150:  # keep the whole run within a few minutes
            break
    dt = (time.time() - t0) / steps
    v = round(1.0 / dt, 4)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": warm, "ms_per_step": round(dt * 1e3, 2), "higher_is_better": True,
            "scaling": "strong" if args.views_per_step > 0 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args.workload, scene, W, H, args, world),
                       "where": "host cores; each step = one view fwd+bwd (a bounded sample of the workload)"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "cpu": cpu_model(),
                             "sample": f"{steps} full iterations, oracle port (torch-CPU skinning + OpenMP C rasterizer)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(line)


def run_reference_sequence(args, scene, world, cores):
    """Forward-only frames of the novel-pose sequence on the host cores (a bounded sample of frames)."""
    import math
    import numpy as np
    import torch
    from oracle import oracle_py as orc
    from oracle import skin_oracle as so
    cam = scene.cameras[0]
    W, H = cam.image_width, cam.image_height
    tfx, tfy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
    view_m, proj_m, campos = cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), cam.camera_center.numpy()
    shs = torch.cat([scene.features_dc, scene.features_rest], 1).numpy()
    bg = np.zeros(3, np.float32)

    def frame(f):
        p = scene.poses[f % len(scene.poses)]
        with torch.no_grad():
            d = so.deform_by_pose(scene.xyz, scene.joints, scene.parents, scene.node_radius_raw, p["local_rotation"],
                                  p["global_trans"], scene.motion_mask)
            m, s, r, o = so.activate(scene.xyz, scene.scaling_raw, scene.rotation_raw, scene.opacity_raw, d["d_xyz"], d["d_rotation"])
        orc.forward(m.numpy(), o.numpy(), view_m, proj_m, campos, W, H, tfx, tfy, bg, shs=shs, scales=s.numpy(), rotations=r.numpy())

    frame(0)
    n, t0 = 0, time.time()
    while n < max(args.steps, 1) * 4 and time.time() - t0 < 120:
        frame(n)
        n += 1
    dt = (time.time() - t0) / n
    v = round(1.0 / dt, 4)
    line = {"impl": "reference", "metric": "novel-pose frames/s (skin + rasterize, forward only)", "value": v, "unit": "frames/s",
            "n_gpus": world, "steps": n, "warmup": 1, "ms_per_step": round(dt * 1e3, 2), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(args.workload, scene, W, H, args, world), "where": "host cores; each step = one frame"},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "cpu": cpu_model(),
                             "sample": f"{n} frames of the sequence, oracle port"},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    _emit(line)


# ---------------------------------------------------------------------------------------------------------
# BASELINE.md rows B1 / B2
def reference_cpu_skinning():
    """B1: the reference's torch CPU skinning path ALONE (skin_oracle.deform_by_pose, the restatement pinned to
    SkeletonWarp.deform_by_pose) on C1 = 100 k Gaussians, J = 24, one frame, forward: all cores and one thread,
    MLPs off and on.  The reference's own class is timed in the build container by profiles/tools/time_reference_skinning.py
    (/root/reference does not travel to the GPU box)."""
    import torch
    from oracle import skin_oracle as so
    from riggs_b200 import synth
    scene = synth.make_config("C1")
    wm, dn = _cpu_mlps(scene)
    p = scene.poses[0]
    cores = os.cpu_count() or 1

    def run(mlps):
        with torch.no_grad():
            mw = off = None
            if mlps:
                mw = wm(scene.xyz)
                off = dn(scene.xyz, p["local_rotation"].reshape(-1)[None].expand(scene.N, -1))
            so.deform_by_pose(scene.xyz, scene.joints, scene.parents, scene.node_radius_raw, p["local_rotation"],
                              p["global_trans"], scene.motion_mask, mw, off)

    def timed(mlps, threads, reps):
        torch.set_num_threads(threads)
        run(mlps)
        ts = []
        for _ in range(reps):
            t0 = time.time()
            run(mlps)
            ts.append(time.time() - t0)
        ts.sort()
        return round(ts[len(ts) // 2] * 1e3, 2)

    out = {"workload": "C1: 100000 Gaussians, 24-joint skeleton, 1 frame, deform_by_pose forward", "cpu": cpu_model(),
           "os_cpu_count": cores, "kind": "port (oracle/skin_oracle.py, pinned to the reference's SkeletonWarp.deform_by_pose)",
           "mlps_off_all_cores_ms": timed(False, cores, 10), "mlps_off_1_thread_ms": timed(False, 1, 3),
           "mlps_on_all_cores_ms": timed(True, cores, 5), "mlps_on_1_thread_ms": timed(True, 1, 1)}
    torch.set_num_threads(cores)
    out["torch_num_threads"] = torch.get_num_threads()
    return out


def reference_cuda_rasterizer(workloads):
    """B2: the reference's CUDA rasterizer ALONE (oracle/_ref, unmodified sources), forward and backward of one view, CUDA
    events on the legacy default stream it launches on, beside this library's reference-shaped entry on the same inputs."""
    try:
        import torch
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import helpers
        from oracle import ref_cuda
        from riggs_b200 import synth
        if not ref_cuda.available():
            return {"unavailable": "oracle/_ref/libref_riggs.so not built"}
        rows = {}
        for name in workloads:
            scene = synth.make_config(name, n_views=1, n_poses=1) if name != "C5" else synth.make_config(name, n_poses=2)
            cin = helpers.to_cuda(helpers.raster_inputs(scene, view=0))
            cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
            ref = ref_cuda.RefRasterizer()

            def fwd():
                return ref.forward(cin["bg"], cin["means3D"], cin["opacities"], cin["view"], cin["proj"], cin["campos"], cin["W"],
                                   cin["H"], cin["tanfovx"], cin["tanfovy"], shs=cin["shs"], scales=cin["scales"], rotations=cin["rotations"])

            def bwd(f):
                return ref.backward(cin["bg"], cin["means3D"], f["radii"], f["alpha"], cin["view"], cin["proj"], cin["campos"],
                                    cin["tanfovx"], cin["tanfovy"], cots[0], cots[1], cots[2], shs=cin["shs"], scales=cin["scales"],
                                    rotations=cin["rotations"])

            def ev_time(fn, reps=5):
                fn()
                ts = []
                for _ in range(reps):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    r = fn()
                    torch.cuda.synchronize()
                    ts.append((time.perf_counter() - t0) * 1e3)
                ts.sort()
                return ts[len(ts) // 2], r

            t_f, f = ev_time(fwd)
            t_b, _ = ev_time(lambda: bwd(f))
            # ours, reference-shaped entry (rg_rasterize_forward / backward through the _C binding), same inputs
            from riggs_b200.diff_gaussian_rasterization import _C
            e = torch.empty(0)

            def ours_fwd():
                return _C.rasterize_gaussians(cin["bg"], cin["means3D"], e, cin["opacities"], cin["scales"], cin["rotations"], 1.0, e,
                                              cin["view"], cin["proj"], cin["tanfovx"], cin["tanfovy"], cin["H"], cin["W"], cin["shs"], 3,
                                              cin["campos"], False, False)

            st = helpers.ours_forward(cin)
            t_of, _ = ev_time(ours_fwd)
            t_ob, _ = ev_time(lambda: helpers.ours_backward(cin, st, cots) and None)
            rows[name] = {"P": scene.N, "R": int(f["num_rendered"]), "W": cin["W"], "H": cin["H"],
                          "reference_forward_ms": round(t_f, 3), "reference_backward_ms": round(t_b, 3),
                          "ours_reference_shaped_forward_ms": round(t_of, 3),
                          "ours_reference_shaped_backward_ms": round(t_ob, 3)}
            del cin, cots, ref, st, f
            torch.cuda.empty_cache()
        rows["how"] = ("median of 5, host timer between device synchronisations (the reference binding allocates its outputs and "
                       "launches on the legacy default stream; its forward includes one blocking 4-byte read-back)")
        return rows
    except Exception as ex:
        return {"unavailable": repr(ex)[:200]}


def ref_cuda_baseline(scene, cam, pose, target, steps=10, mlps=None):
    """Context number asked by the north star: the reference's own CUDA rasterizer (compiled unmodified,
    oracle/_ref) behind an eager-torch skinning chain on the same B200, with the share of each stage.  Not the reference arm."""
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
        stage = {"skinning_forward": 0.0, "rasterizer_forward": 0.0, "rasterizer_backward": 0.0, "skinning_backward": 0.0}

        def lap(key, t0):
            torch.cuda.synchronize()
            stage[key] += time.perf_counter() - t0
            return time.perf_counter()

        def step(record):
            t0 = time.perf_counter()
            q = pose["local_rotation"].detach().clone().requires_grad_(True)
            g = pose["global_trans"].detach().clone().requires_grad_(True)
            mw = off = None
            if mlps is not None:  # the same torch modules the GPU arm's reference would run eagerly (cuBLAS)
                mw = mlps.weight_mlp(xyz.detach())
                off = mlps.detail_net(xyz.detach(), q.detach().reshape(-1)[None].expand(xyz.shape[0], -1))
            d = so.deform_by_pose(xyz, joints, parents, nr, q, g, mask, mw, off)
            means3D, scales, rots, opac = so.activate(xyz, sc_raw, rot_raw, op_raw, d["d_xyz"], d["d_rotation"])
            shs = torch.cat([fdc, frest], 1)
            a = [t.detach().contiguous() for t in (means3D, opac, shs, scales, rots)]
            if record:
                t0 = lap("skinning_forward", t0)
            f = ref.forward(bg, a[0], a[1], view_m, proj_m, campos, W, H, tfx, tfy, shs=a[2], scales=a[3], rotations=a[4])
            if record:
                t0 = lap("rasterizer_forward", t0)
            gimg = (torch.sign(f["color"] - target) / target.numel()).contiguous()
            z = torch.zeros(1, H, W, device=dev)
            gr = ref.backward(bg, a[0], f["radii"], f["alpha"], view_m, proj_m, campos, tfx, tfy, gimg, z, z, shs=a[2],
                              scales=a[3], rotations=a[4])
            if record:
                t0 = lap("rasterizer_backward", t0)
            torch.autograd.backward([means3D, scales, rots, opac, shs],
                                    [gr["dmean3D"], gr["dscale"], gr["drot"], gr["dopacity"], gr["dsh"]])
            for t in (xyz, sc_raw, rot_raw, op_raw, fdc, frest, nr):
                t.grad = None
            if record:
                lap("skinning_backward", t0)

        for _ in range(3):
            step(False)
        torch.cuda.synchronize()
        t0 = time.time()
        for _ in range(steps):
            step(False)
        torch.cuda.synchronize()
        dt = (time.time() - t0) / steps
        for _ in range(3):
            step(True)
        return {"value": round(1.0 / dt, 3), "unit": UNIT, "ms_per_step": round(dt * 1e3, 3),
                "stages_ms": {k: round(v / 3 * 1e3, 3) for k, v in stage.items()},
                "what": "eager-torch skinning on CUDA + the reference's CUDA rasterizer (oracle/_ref, unmodified sources) fwd+bwd"
                        + (", both MLPs eager torch/cuBLAS" if mlps is not None else "")}
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
