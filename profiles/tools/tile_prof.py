"""Per-warp cycle counters of the blend kernels on the C0 frame (rg_debug_tile_profile): how long the slowest warps run
against the kernel's duration, and what a warp's cycles go to."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from riggs_b200 import synth, _lib
from riggs_b200.hotpath import HotPath
scene = synth.make_config("C0", n_views=8)
hp = HotPath(scene, device="cuda")
cam = scene.cameras[0].to("cuda")
pose = {k: v.cuda().requires_grad_(True) for k, v in scene.poses[0].items()}
target = torch.rand(3, 800, 800, device="cuda")
L = _lib.lib()
T = 2500
for _ in range(3):
    hp.zero_grad(); hp.step(cam, pose, target)
fw = torch.zeros(T * 8 * 8, dtype=torch.int32, device="cuda"); bw = torch.zeros_like(fw)
L.rg_debug_tile_profile(fw.data_ptr(), bw.data_ptr())
hp.zero_grad(); hp.step(cam, pose, target); torch.cuda.synchronize()
L.rg_debug_tile_profile(None, None)
ghz = 1.965
for name, t, slots in (("fwd", fw, 148 * 16), ("bwd", bw, 148 * 16)):
    a = t.cpu().numpy().astype(np.int64).reshape(T * 8, 8)
    tot = a[:, 0]
    live = tot > 0
    srt = np.sort(tot)[::-1]
    print(f"{name}: warps {live.sum()}, sum of warp cycles {tot.sum()} -> balanced over {slots} slots {tot.sum() / slots / ghz / 1e3:.1f} us; "
          f"slowest warp {srt[0] / ghz / 1e3:.1f} us, 10th {srt[9] / ghz / 1e3:.1f}, 100th {srt[99] / ghz / 1e3:.1f}, "
          f"p99 {np.percentile(tot[live], 99) / ghz / 1e3:.1f}, median {np.median(tot[live]) / ghz / 1e3:.2f} us")
    print(f"   shares of all warp cycles: wait {a[:, 1].sum() / tot.sum():.3f} stage1 {a[:, 2].sum() / tot.sum():.3f} stage2 {a[:, 3].sum() / tot.sum():.3f}; "
          f"survivors {a[:, 4].sum()}, list entries walked {a[:, 6].sum() if name == 'fwd' else 0}")
    top = np.argsort(-tot)[:5]
    for i in top:
        print("   warp", i // 8, i % 8, "cycles", a[i, 0], "wait/stage1/stage2", a[i, 1], a[i, 2], a[i, 3], "survivors", a[i, 4], "groups", a[i, 5], "aux", a[i, 6])
