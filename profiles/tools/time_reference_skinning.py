"""BASELINE.md row B1, measured on the REFERENCE's own code: SkeletonWarp.deform_by_pose / .forward of /root/reference
(skeleton_utils/skeleton_warp.py:112-172) on C1 = 100 k Gaussians, 24 joints, one frame, torch on the CPU.

Runs in the BUILD CONTAINER only (the reference tree does not travel to the GPU box; there bench.py --baselines times
oracle/skin_oracle.py, the restatement pinned to this class).  Import shims as in SURVEY.md section 8(c).
    python profiles/tools/time_reference_skinning.py > profiles/r2_reference_cpu_skinning_container.json
"""
import json
import os
import sys
import time
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
p3d, ops = types.ModuleType("pytorch3d"), types.ModuleType("pytorch3d.ops")
ops.knn_points = ops.ball_query = None
p3d.ops = ops
sys.modules["pytorch3d"], sys.modules["pytorch3d.ops"] = p3d, ops
torch.Tensor.cuda = lambda self, *a, **k: self
torch.nn.Module.cuda = lambda self, *a, **k: self

import contextlib
import io

with contextlib.redirect_stdout(io.StringIO()):
    from skeleton_utils.skeleton_warp import SkeletonWarp  # noqa: E402
from riggs_b200 import synth  # noqa: E402


def cpu_model():
    for line in open("/proc/cpuinfo"):
        if line.startswith("model name"):
            return line.split(":", 1)[1].strip()
    return "unknown"


def main():
    sc = synth.make_config("C1")
    with contextlib.redirect_stdout(io.StringIO()):
        m = SkeletonWarp(is_blender=True, joints=sc.joints, parent_indices=sc.parents, K=-1,
                         use_skinning_weight_mlp=True, use_template_offsets=True)
    m._node_radius.data = sc.node_radius_raw.clone()
    pose = {"local_rotation": sc.poses[0]["local_rotation"], "global_trans": sc.poses[0]["global_trans"]}
    t = torch.tensor([[0.3]]).expand(sc.J, 1)
    cores = os.cpu_count() or 1

    def timed(fn, threads, reps):
        torch.set_num_threads(threads)
        with torch.no_grad():
            fn()
            ts = []
            for _ in range(reps):
                t0 = time.time()
                fn()
                ts.append(time.time() - t0)
        ts.sort()
        return round(ts[len(ts) // 2] * 1e3, 2)

    out = {"what": "the reference's own SkeletonWarp on the CPU of the build container (not the GPU box)",
           "workload": "C1: 100000 Gaussians, 24-joint skeleton, 1 frame, forward", "cpu": cpu_model(), "os_cpu_count": cores}
    for mlps in (False, True):
        m.use_skinning_weight_mlp = mlps
        m.use_template_offsets = mlps
        tag = "mlps_on" if mlps else "mlps_off"
        out[f"deform_by_pose_{tag}_all_cores_ms"] = timed(lambda: m.deform_by_pose(sc.xyz, pose, sc.motion_mask), cores, 5 if mlps else 10)
        out[f"deform_by_pose_{tag}_1_thread_ms"] = timed(lambda: m.deform_by_pose(sc.xyz, pose, sc.motion_mask), 1, 1 if mlps else 3)
        out[f"forward_incl_pose_mlp_{tag}_all_cores_ms"] = timed(lambda: m.forward(sc.xyz, t, sc.motion_mask), cores, 5 if mlps else 10)
    torch.set_num_threads(cores)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
