import sys, math, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import helpers
from riggs_b200 import synth
from riggs_b200.fused import render_skinned, WorkspacePool
from riggs_b200.diff_gaussian_rasterization import _C
import test_gpu_fused as T
scene, pc, warp, cam = T._setup("small", {})
P, W, H = scene.N, cam.image_width, cam.image_height
bg = torch.tensor([0.1, 0.3, 0.2], device="cuda")
pose = {k: v.cuda() for k, v in scene.poses[0].items()}
dbg = dict(scales=torch.empty(P, 3, device="cuda"), rotations=torch.empty(P, 4, device="cuda"), opacities=torch.empty(P, 1, device="cuda"))
pool = WorkspacePool()
with torch.no_grad():
    out = render_skinned(cam, pc, warp, pose, T._Pipe(), bg, debug_outputs=dbg, pool=pool)
torch.cuda.synchronize()
ws = pool.free[(P, W, H, str(pc._xyz.device))][0]
R = int(ws.status[0])
ptrs = _C.state_ptrs(P, R, W, H, ws.geom, ws.binning, ws.image)
means_f = helpers._slice(ws.geom, ptrs[7], 12 * P, torch.float32).reshape(P, 3).clone()
def geomarrays(geom, ptrs):
    return dict(depths=helpers._slice(geom, ptrs[0], 4*P, torch.float32).clone(), means2D=helpers._slice(geom, ptrs[2], 8*P, torch.float32).clone(),
                conic=helpers._slice(geom, ptrs[3], 16*P, torch.float32).clone(), rgb=None)
A = geomarrays(ws.geom, ptrs)
shs = torch.cat([pc._features_dc, pc._features_rest], 1).detach().contiguous()
e = torch.empty(0)
R2, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(bg, means_f, e, dbg["opacities"], dbg["scales"], dbg["rotations"], 1.0, e, cam.world_view_transform, cam.full_proj_transform, math.tan(cam.FoVx*0.5), math.tan(cam.FoVy*0.5), H, W, shs, 3, cam.camera_center, False, False)
p2 = _C.state_ptrs(P, R2, W, H, geom, binning, img)
B = geomarrays(geom, p2)
vis = (radii > 0)
for k in A:
    a, b = A[k].reshape(P, -1)[vis], B[k].reshape(P, -1)[vis]
    nd = (a != b).any(1).sum().item()
    print(k, 'rows differing', nd, 'max rel', ((a-b).abs().max() / b.abs().max()).item())
print('color equal', torch.equal(color, out['render']), (color-out['render']).abs().max().item())
