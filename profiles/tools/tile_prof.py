import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
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
fw = torch.zeros(T*8*6, dtype=torch.int32, device="cuda"); bw = torch.zeros_like(fw)
L.rg_debug_tile_profile(fw.data_ptr(), bw.data_ptr())
hp.zero_grad(); hp.step(cam, pose, target); torch.cuda.synchronize()
L.rg_debug_tile_profile(None, None)
for name, t in (("fwd", fw), ("bwd", bw)):
    a = t.cpu().numpy().astype(np.int64).reshape(T, 8, 6)
    tot = a[:, :, 0].max(1)   # per tile: slowest warp total
    order = np.argsort(-tot)
    print(name, "sum over tiles of max-warp cycles", tot.sum(), "max tile", tot.max(), "mean nonzero", tot[tot>0].mean(), "n>0", (tot>0).sum())
    print("  ideal balanced (sum/148/occupancy~4):", tot.sum()/148/4)
    for t_ in order[:6]:
        w = a[t_]
        print("  tile", t_, "total", w[:,0].max(), "filter(max/mean)", w[:,1].max(), int(w[:,1].mean()), "list(max/mean)", w[:,2].max(), int(w[:,2].mean()), "entries(max/mean)", w[:,3].max(), int(w[:,3].mean()), "batches", w[:,4].max(), "aux", w[:,5].max())
    # aggregate cycle shares
    print("  share filter", a[:,:,1].sum()/a[:,:,0].sum(), "list", a[:,:,2].sum()/a[:,:,0].sum(), "entries total", a[:,:,3].sum(), "cycles/list entry", a[:,:,2].sum()/max(a[:,:,3].sum(),1), "filter cycles/batch", a[:,:,1].sum()/max(a[:,:,4].sum(),1))
    hist = np.percentile(tot[tot>0], [50, 90, 99, 100]); print("  tile total cycles p50/p90/p99/max", hist)
