"""GPU parity at the sizes BASELINE.json names: C0 (300 k Gaussians, J = 24, 800 x 800, the headline configuration) and
C2 (200 k, J = 20, 800 x 800), forward AND backward, through the reference-shaped C-ABI entry against the reference's own
CUDA kernels (oracle/_ref).  The heaviest 8x4 block of C0 walks ~3 500 list entries: the paths only long lists reach
(more than 32 survivors per group, hundreds of chunk flushes per warp, suffix sums over thousands of terms) are gated
here for gradients, not only for images.

Gates: forward integers bit-exact; images 1e-5; every gradient array 1e-4 relative to its largest entry (the north
star's statement) AND 1e-4 in relative L2 (pins the bulk of the entries, which a max-norm ratio lets through) AND a
per-entry comparison with a floor at 1e-3 of the largest entry.  The reference accumulates with float atomics in an
order that changes from run to run; the per-entry gate is therefore stated against the reference's own run-to-run
spread (measured in the same test) with an absolute ceiling."""
import json
import os

import numpy as np
import pytest
import torch

import helpers
from oracle import ref_cuda
from riggs_b200 import synth
from test_gpu_raster import _run_ref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref/libref_riggs.so not built")]

GRADS = ("dmean2D", "dcolor", "dopacity", "dmean3D", "dcov3D", "dsh", "dscale", "drot")
# Per-entry statistic (floor at 1e-3 of the largest entry).  Measured at C0 (gpurun_out -> profiles/r2_parity_fullsize.md):
# the reference against a second run of itself 3e-5 .. 1.4e-3 (drot), this library against the reference 9e-5 .. 2.7e-3,
# i.e. 1.2x .. 4.7x the reference's own run-to-run spread (a maximum over 10^5 .. 10^7 entries, itself a noisy figure).
FLOORED_CEIL = 5e-3
FLOORED_VS_REF_SPREAD = 8.0


def _record(name, stats):
    out = os.path.join(helpers.ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, f"parity_{name}.json"), "w") as f:
            json.dump(stats, f, indent=1)
    except OSError:
        pass


@pytest.mark.parametrize("name", ["C0", "C2"])
def test_full_size_forward_backward_vs_reference_cuda(name):
    scene = synth.make_config(name, n_views=1)
    cin = helpers.to_cuda(helpers.raster_inputs(scene, view=0))
    cots = [t.cuda() for t in synth.seeded_cotangents(cin["H"], cin["W"])]
    ours = helpers.ours_forward(cin)
    g = helpers.ours_backward(cin, ours, cots)
    f, st, gr = _run_ref(cin, cots)
    _, _, gr2 = _run_ref(cin, cots)  # the reference against itself: its atomics' run-to-run spread
    assert ours["num_rendered"] == f["num_rendered"]
    assert np.array_equal(ours["radii"], f["radii"].cpu().numpy())
    assert np.array_equal(ours["tiles_touched"], st["tiles_touched"])
    vis = ours["radii"] > 0
    assert np.array_equal(ours["depths"][vis].view(np.uint32), st["depths"][vis].view(np.uint32))
    assert np.array_equal(ours["means2D"][vis].view(np.uint32), st["means2D"][vis].view(np.uint32))
    assert np.array_equal(ours["keys"], st["keys"])
    assert np.array_equal(ours["point_list"], st["point_list"])
    assert np.array_equal(ours["ranges"], st["ranges"])
    assert np.array_equal(ours["n_contrib"], st["n_contrib"])
    for k in ("color", "depth", "alpha"):
        assert helpers.rel_err(ours[k].cpu().numpy(), f[k].cpu().numpy()) <= 1e-5, k
    stats = {"R": int(ours["num_rendered"]), "longest_list": int((st["ranges"][:, 1] - st["ranges"][:, 0]).max())}
    for k in GRADS:
        a, b, b2 = (x[k].cpu().numpy().reshape(-1) for x in (g, gr, gr2))
        s = helpers.grad_stats(a, b)
        s["ref_vs_ref"] = helpers.grad_stats(b2, b)
        stats[k] = s
    _record(f"fullsize_{name}", stats)
    for k in GRADS:
        s = stats[k]
        assert s["max"] <= 1e-4, (k, s)
        assert s["l2"] <= 1e-4, (k, s)
        assert s["floored"] <= max(1e-4, min(FLOORED_CEIL, FLOORED_VS_REF_SPREAD * s["ref_vs_ref"]["floored"])), (k, s)
