"""Distribution of the per-8x4-block work of the blend kernels on the C0 frame, from the CPU oracle (no GPU)."""
import math, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import oracle_py as orc, skin_oracle as so
from riggs_b200 import synth
scene = synth.make_config("C0")
cam = scene.cameras[0]; p = scene.poses[0]
d = so.deform_by_pose(scene.xyz, scene.joints, scene.parents, scene.node_radius_raw, p["local_rotation"], p["global_trans"], scene.motion_mask)
means3D, scales, rots, opac = so.activate(scene.xyz, scene.scaling_raw, scene.rotation_raw, scene.opacity_raw, d["d_xyz"], d["d_rotation"])
shs = torch.cat([scene.features_dc, scene.features_rest], 1)
W, H = cam.image_width, cam.image_height
st = orc.forward(means3D.numpy(), opac.numpy(), cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(), cam.camera_center.numpy(),
                 W, H, math.tan(cam.FoVx*.5), math.tan(cam.FoVy*.5), np.zeros(3, np.float32), shs=shs.numpy(), scales=scales.numpy(), rotations=rots.numpy())
print({k: getattr(v, 'shape', v) for k, v in st.items()})
rng = st["ranges"].reshape(-1, 2); ln = (rng[:, 1] - rng[:, 0]).astype(np.int64)
nc = st["n_contrib"].reshape(H, W)
blk = nc.reshape(H // 4, 4, W // 8, 8).max(axis=(1, 3))   # walk length per 8x4 block
print("tiles", ln.size, "list len mean/max", ln.mean(), ln.max(), "p99", np.percentile(ln, 99))
print("blocks", blk.size, "walk len mean/max", blk.mean(), blk.max(), "p99", np.percentile(blk, 99), "sum", blk.sum())
w = np.sort(blk.reshape(-1))[::-1]
slots = 148 * 16
print("sum/slots", w.sum() / slots, "max", w[0], "-> critical path share", w[0] / (w.sum() / slots))
print("top 10 blocks", w[:10])
