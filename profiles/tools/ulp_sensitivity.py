import sys, math, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import helpers
from riggs_b200 import synth
from riggs_b200.diff_gaussian_rasterization import _C
from oracle import ref_cuda
scene = synth.make_config("small")
cin = helpers.to_cuda(helpers.raster_inputs(scene, view=0, bg=(0.1,0.3,0.2)))
cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
def run(c):
    st = helpers.ours_forward(c)
    g = helpers.ours_backward(c, st, cots)
    return st, {k: v.clone() for k, v in g.items()}
st1, g1 = run(cin)
st2, g2 = run(cin)
for k in g1: print('rerun', k, helpers.rel_err(g1[k].cpu().numpy().reshape(-1), g2[k].cpu().numpy().reshape(-1)))
# perturb rotations by 1 ulp
c2 = dict(cin); r = cin["rotations"].clone(); r = torch.nextafter(r, torch.full_like(r, 10.0)); c2["rotations"] = r
st3, g3 = run(c2)
print('img diff under 1ulp rot', helpers.rel_err(st3["color"].cpu().numpy(), st1["color"].cpu().numpy()), 'n_contrib diff', (st3["n_contrib"]!=st1["n_contrib"]).sum())
for k in g1: print('1ulp rot', k, helpers.rel_err(g3[k].cpu().numpy().reshape(-1), g1[k].cpu().numpy().reshape(-1)))
d = (g3["dmean3D"]-g1["dmean3D"]).abs().max(1).values
i = int(d.argmax()); print('worst row', i, g1["dmean3D"][i].tolist(), g3["dmean3D"][i].tolist(), 'radius', st1["radii"][i], 'conic', st1["conic_opacity"][i], st3["conic_opacity"][i], 'dmean2D', g1["dmean2D"][i].tolist(), g3["dmean2D"][i].tolist(), 'dcov', g1["dcov3D"][i].tolist(), g3["dcov3D"][i].tolist())
# reference CUDA same experiment
ref = ref_cuda.RefRasterizer()
def runref(c):
    f = ref.forward(c["bg"], c["means3D"], c["opacities"], c["view"], c["proj"], c["campos"], c["W"], c["H"], c["tanfovx"], c["tanfovy"], shs=c["shs"], scales=c["scales"], rotations=c["rotations"])
    g = ref.backward(c["bg"], c["means3D"], f["radii"], f["alpha"], c["view"], c["proj"], c["campos"], c["tanfovx"], c["tanfovy"], cots[0], cots[1], cots[2], shs=c["shs"], scales=c["scales"], rotations=c["rotations"])
    return {k: v.clone() for k, v in g.items()}
r1 = runref(cin); r1b = runref(cin); r3 = runref(c2)
for k in ("dmean3D","dscale","drot","dmean2D"): print('REF rerun', k, helpers.rel_err(r1b[k].cpu().numpy().reshape(-1), r1[k].cpu().numpy().reshape(-1)), ' REF 1ulp', helpers.rel_err(r3[k].cpu().numpy().reshape(-1), r1[k].cpu().numpy().reshape(-1)))
