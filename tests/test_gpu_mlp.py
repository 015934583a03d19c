"""GPU parity of the tensor-core MLP front end (csrc/mlp.cu, riggs_b200.mlp.FusedSkipMLP) against the torch modules of
skeleton_warp -- which tests/test_mlp_front_end.py pins to the reference's own network_utils.py at 1e-6.

Tolerance: the kernels multiply in bf16 with fp32 accumulation where the reference runs fp32 cuBLAS; eight 256-wide
layers deep that is ~1e-2 of the output's scale at worst (measured and recorded by the test), stated in DESIGN.md."""
import json
import os

import numpy as np
import pytest
import torch

import helpers
from riggs_b200 import skeleton_warp as sw

pytestmark = pytest.mark.gpu

FWD_TOL = 2e-2   # max |fused - fp32 torch| / max |fp32 torch|


def _modules(J, seed=3):
    torch.manual_seed(seed)
    wm = sw.WeightMLP(3, J - 1).cuda()
    dn = sw.DeformMLP(4 * J).cuda()
    with torch.no_grad():
        dn.gaussian_warp.weight.normal_(0, 0.05)
        dn.gaussian_warp.bias.normal_(0, 0.01)
    return wm, dn


@pytest.mark.parametrize("N,J", [(128, 24), (1000, 24), (40_000, 60), (300, 7)])
def test_fused_mlp_forward_matches_torch(N, J):
    from riggs_b200.mlp import FusedSkipMLP
    wm, dn = _modules(J)
    g = torch.Generator().manual_seed(N)
    x = (torch.randn(N, 3, generator=g) * 0.6).cuda()
    pose = (torch.randn(4 * J, generator=g) * 0.3).cuda()
    fw, fd = FusedSkipMLP(wm), FusedSkipMLP(dn)
    with torch.no_grad():
        want_w = wm(x)
        want_d = dn(x, pose[None].expand(N, -1))
    got_w = fw(x)
    got_d = fd(x, pose)
    torch.cuda.synchronize()
    ew = helpers.rel_err(got_w.cpu().numpy(), want_w.cpu().numpy())
    ed = helpers.rel_err(got_d.cpu().numpy(), want_d.cpu().numpy())
    try:
        with open(os.path.join(helpers.ROOT, "gpurun_out", f"mlp_fwd_N{N}_J{J}.json"), "w") as f:
            json.dump({"weight_mlp_rel_err": ew, "detail_net_rel_err": ed,
                       "weight_mlp_l2": helpers.rel_l2(got_w.cpu().numpy(), want_w.cpu().numpy()),
                       "detail_net_l2": helpers.rel_l2(got_d.cpu().numpy(), want_d.cpu().numpy())}, f)
    except OSError:
        pass
    assert got_w.shape == want_w.shape and got_d.shape == want_d.shape
    assert torch.isfinite(got_w).all() and torch.isfinite(got_d).all()
    assert ew <= FWD_TOL, ew
    assert ed <= FWD_TOL, ed
    # a second call (weights unchanged: no repack) gives the same bits
    assert torch.equal(fw(x), got_w)
