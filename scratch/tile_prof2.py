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
fw = torch.zeros(T*8*8, dtype=torch.int32, device="cuda"); bw = torch.zeros_like(fw)
L.rg_debug_tile_profile(fw.data_ptr(), bw.data_ptr())
hp.zero_grad(); hp.step(cam, pose, target); torch.cuda.synchronize()
L.rg_debug_tile_profile(None, None)
for name, t in (("fwd", fw), ("bwd", bw)):
    a = t.cpu().numpy().astype(np.int64).reshape(T*8, 8)
    cyc, tw, t1, t2, nl, nb, aux = [a[:, i] for i in range(7)]
    act = cyc > 0
    print(name, "warps active", act.sum(), "sum cycles", cyc.sum(), "max", cyc.max(), "p50/p90/p99", np.percentile(cyc[act],[50,90,99]))
    print("   sum groups", nb.sum(), "sum list", nl.sum(), "aux sum", aux.sum())
    print("   shares: wait %.3f stage1 %.3f stage2 %.3f rest(list loop) %.3f" % (tw.sum()/cyc.sum(), t1.sum()/cyc.sum(), t2.sum()/cyc.sum(), 1-(tw.sum()+t1.sum()+t2.sum())/cyc.sum()))
    print("   per group: wait %.0f stage1 %.0f stage2 %.0f ; list loop per entry %.0f" % (tw.sum()/nb.sum(), t1.sum()/nb.sum(), t2.sum()/nb.sum(), (cyc.sum()-tw.sum()-t1.sum()-t2.sum())/max(nl.sum(),1)))
    top = np.argsort(-cyc)[:6]
    for i in top: print("   warp", i//8, i%8, a[i])
