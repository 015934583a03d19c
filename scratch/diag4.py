import sys, math, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import helpers
from riggs_b200 import synth
from riggs_b200.fused import render_skinned, WorkspacePool
from riggs_b200.diff_gaussian_rasterization import _C
from oracle import ref_cuda
import test_gpu_fused as T
for over in (dict(isotropic=True, soft_mask=True), dict(soft_mask=True, with_mlp_weight=True, with_offsets=True)):
    scene, pc, warp, cam = T._setup("small", over)
    P, W, H = scene.N, cam.image_width, cam.image_height
    bg = torch.zeros(3, device="cuda")
    pose = {k: v.cuda() for k, v in scene.poses[0].items()}
    dbg = dict(scales=torch.empty(P, 3, device="cuda"), rotations=torch.empty(P, 4, device="cuda"), opacities=torch.empty(P, 1, device="cuda"))
    pool = WorkspacePool()
    with torch.no_grad():
        out = render_skinned(cam, pc, warp, pose, T._Pipe(), bg, debug_outputs=dbg, pool=pool,
              mlp_weight=scene.mlp_weight.cuda() if scene.mlp_weight is not None else None, offsets=scene.template_offsets.cuda() if scene.template_offsets is not None else None)
    torch.cuda.synchronize()
    ws = pool.free[(P, W, H, str(pc._xyz.device))][0]
    R = int(ws.status[0])
    ptrs = _C.state_ptrs(P, R, W, H, ws.geom, ws.binning, ws.image)
    means_f = helpers._slice(ws.geom, ptrs[7], 12 * P, torch.float32).reshape(P, 3).clone()
    shs = torch.cat([pc._features_dc, pc._features_rest], 1).detach().contiguous()
    ref = ref_cuda.RefRasterizer()
    f = ref.forward(bg, means_f, dbg["opacities"], cam.world_view_transform.contiguous(), cam.full_proj_transform.contiguous(), cam.camera_center.contiguous(), W, H, math.tan(cam.FoVx*0.5), math.tan(cam.FoVy*0.5), shs=shs, scales=dbg["scales"], rotations=dbg["rotations"])
    bad = (out["radii"] != f["radii"]).nonzero().flatten()
    print(over, 'radii mismatches', bad.numel(), 'R', R, f["num_rendered"])
    for i in bad[:5].tolist():
        print(i, out["radii"][i].item(), f["radii"][i].item(), dbg["scales"][i].tolist(), dbg["rotations"][i].tolist())
    # unfused ours on same inputs
    e = torch.empty(0)
    R2, color, depth, alpha, radii, geom, binning, img = _C.rasterize_gaussians(bg, means_f, e, dbg["opacities"], dbg["scales"], dbg["rotations"], 1.0, e, cam.world_view_transform, cam.full_proj_transform, math.tan(cam.FoVx*0.5), math.tan(cam.FoVy*0.5), H, W, shs, 3, cam.camera_center, False, False)
    print('unfused ours vs ref radii mismatches', (radii != f["radii"]).sum().item())
