"""Stand-in for ``simple_knn._C`` (knn/ext.cpp:16, knn/spatial.cu:15-26)."""
from __future__ import annotations

import torch

from .. import _lib
from .._lib import check, f32c, ptr, stream_ptr


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """Mean squared distance of each point to its 3 nearest neighbours; (P,3) fp32 CUDA -> (P,) fp32."""
    if not points.is_cuda:
        raise _lib.RiggsError("distCUDA2 needs a CUDA tensor; riggs_b200 has no CPU path")
    P = points.size(0)
    means = torch.zeros((P,), dtype=torch.float32, device=points.device)
    if P == 0:
        return means
    pts = f32c(points)
    nbytes = _lib.lib().rg_knn_scratch_bytes(P)
    scratch = torch.empty(int(nbytes), dtype=torch.uint8, device=points.device)
    with torch.cuda.device(points.device):
        check(_lib.lib().rg_dist_cuda2(P, ptr(pts), ptr(means), ptr(scratch), nbytes, stream_ptr()), "distCUDA2")
    return means
