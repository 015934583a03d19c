import sys, math, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import helpers
from riggs_b200 import synth, skeleton_warp as sw
from riggs_b200.fused import render_skinned, WorkspacePool
from riggs_b200.gaussian_renderer import render, GaussianParams
from riggs_b200.diff_gaussian_rasterization import _C
import test_gpu_fused as T
scene, pc, warp, cam = T._setup("small", {})
bg = torch.tensor([0.1, 0.3, 0.2], device="cuda")
cots = [t.cuda() for t in synth.seeded_cotangents(cam.image_height, cam.image_width)]
pose, mlp, off = T._leafs(scene)
joints = warp.nodes[:, :3]
G, node_rot, d_nodes = sw.fk(pose["local_rotation"], pose["global_trans"], joints, warp.parents_i32)
d_xyz, d_rot, _ = sw.skin(pc.get_xyz, joints, warp.parents_i32, warp._node_radius, G, node_rot, pose["global_trans"], pc._motion_mask.reshape(-1), None, None)
out_u = render(cam, pc, T._Pipe(), bg, d_xyz, d_rot, 0.0)
T._loss(out_u, cots).backward()
gu = pc._xyz.grad.clone(); pc._xyz.grad=None
means_u = (pc.get_xyz + d_xyz).detach()
rot_u = pc.get_rotation_bias(d_rot).detach()
for p in [pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest, warp._node_radius]: p.grad=None
pose, mlp, off = T._leafs(scene)
P=scene.N
dbg = dict(scales=torch.empty(P, 3, device="cuda"), rotations=torch.empty(P, 4, device="cuda"), opacities=torch.empty(P, 1, device="cuda"))
pool=WorkspacePool()
out_f = render_skinned(cam, pc, warp, pose, T._Pipe(), bg, debug_outputs=dbg, pool=pool)
T._loss(out_f, cots).backward()
gf = pc._xyz.grad.clone()
torch.cuda.synchronize()
ws = pool.free[(P, cam.image_width, cam.image_height, str(pc._xyz.device))][0]
ptrs = _C.state_ptrs(P, int(ws.status[0]), cam.image_width, cam.image_height, ws.geom, ws.binning, ws.image)
means_f = helpers._slice(ws.geom, ptrs[7], 12 * P, torch.float32).reshape(P, 3).clone()
print('means3D max abs diff', (means_f-means_u).abs().max().item(), 'n diff', (means_f!=means_u).any(1).sum().item())
print('rot max abs diff', (dbg['rotations']-rot_u).abs().max().item())
print('scale diff', (dbg['scales']-pc.get_scaling).abs().max().item(), 'opac', (dbg['opacities']-pc.get_opacity).abs().max().item())
d=(gf-gu).abs(); print('grad max', gu.abs().max().item(), 'max diff', d.max().item())
idx=d.max(1).values.topk(8).indices
for i in idx.tolist():
    print(i, gf[i].tolist(), gu[i].tolist(), 'radii', out_f['radii'][i].item(), out_u['radii'][i].item(), 'means diff', (means_f[i]-means_u[i]).tolist())
print('img diff', (out_f['render']-out_u['render']).abs().max().item())
