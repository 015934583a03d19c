"""Capture outputs of the REFERENCE CUDA rasterizer (oracle/_ref/libref_riggs.so = the reference's own
.cu files compiled unmodified) on seeded synthetic inputs, as fixtures for the CPU-only test-suite.

Run on a B200 box:   gpurun -- python tests/golden/make_ref_cuda_golden.py
writes gpurun_out/ref_cuda_<case>.npz; copy them to tests/golden/.
Inputs are NOT stored: they are regenerated from riggs_b200.synth with the recorded config name.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
from oracle import ref_cuda  # noqa: E402
from riggs_b200 import synth  # noqa: E402

CASES = {"ragged": dict(config="ragged", bg=(0.2, 0.4, 0.6)), "mini": dict(config="tiny", bg=(0.0, 0.0, 0.0), N=700)}


def main():
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    for name, c in CASES.items():
        over = {k: v for k, v in c.items() if k not in ("config", "bg")}
        scene = synth.make_config(c["config"], **over)
        inp = helpers.raster_inputs(scene, view=0, bg=c["bg"])
        cin = helpers.to_cuda(inp)
        ref = ref_cuda.RefRasterizer()
        f = ref.forward(cin["bg"], cin["means3D"], cin["opacities"], cin["view"], cin["proj"], cin["campos"], cin["W"],
                        cin["H"], cin["tanfovx"], cin["tanfovy"], shs=cin["shs"], scales=cin["scales"],
                        rotations=cin["rotations"])
        st = ref.state_arrays()
        cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
        g = ref.backward(cin["bg"], cin["means3D"], f["radii"], f["alpha"], cin["view"], cin["proj"], cin["campos"],
                         cin["tanfovx"], cin["tanfovy"], cots[0], cots[1], cots[2], shs=cin["shs"],
                         scales=cin["scales"], rotations=cin["rotations"])
        vis = f["radii"].cpu().numpy() > 0
        d = dict(num_rendered=np.int64(f["num_rendered"]), radii=f["radii"].cpu().numpy(),
                 color=f["color"].cpu().numpy(), depth=f["depth"].cpu().numpy(), alpha=f["alpha"].cpu().numpy(),
                 tiles_touched=st["tiles_touched"], visible=vis, n_contrib=st["n_contrib"], ranges=st["ranges"],
                 keys=st.get("keys", np.zeros(0, np.uint64)), point_list=st.get("point_list", np.zeros(0, np.uint32)))
        # per-Gaussian floats are only defined for visible Gaussians (the reference leaves the rest unwritten)
        for k in ("depths", "means2D", "conic_opacity", "rgb", "clamped"):
            d[k] = np.where(vis.reshape(-1, *([1] * (st[k].ndim - 1))), st[k], 0)
        for k, v in g.items():
            d["g_" + k] = v.cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, f"ref_cuda_{name}.npz"), **d)
        print(name, "R", f["num_rendered"], "visible", int(vis.sum()))


if __name__ == "__main__":
    main()
