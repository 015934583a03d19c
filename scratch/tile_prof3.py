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
a = bw.cpu().numpy().astype(np.int64).reshape(T*8, 8)
act = a[:,0] > 0
s = a[act].sum(0)
print("bwd: total", s[0], "walk", s[1]+s[2]+s[3], "entries", s[4], "A(indep)", s[5], "B(serial)", s[6], "C(park+flush+loop)", s[7])
print("per entry: A %.0f B %.0f C %.0f other %.0f" % (s[5]/s[4], s[6]/s[4], s[7]/s[4], (s[0]-s[1]-s[2]-s[3]-s[5]-s[6]-s[7])/s[4]))
top = np.argsort(-a[:,0])[:3]
for i in top: print(a[i])
