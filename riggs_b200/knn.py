"""``knn_points`` with the call signature the reference uses from pytorch3d.ops (utils/time_utils.py:950, :980, :1051,
:1061, :1124), on the brute-force kernel of csrc/knn_points.cu.  pytorch3d is a git-main dependency of the reference and
is not part of this repository; what is mirrored is its documented contract: squared Euclidean distances of the K
nearest points of p2 for every point of p1, ascending, with their indices (int64) and optionally the neighbours
themselves.  Distances are differentiable (recomputed from the indices with torch when an input requires grad).
No CPU path.
"""
from __future__ import annotations

from collections import namedtuple
from typing import Optional

import torch

from . import _lib
from ._lib import check, f32c, ptr, stream_ptr

_KNN = namedtuple("KNN", "dists idx knn")


def knn_points(p1: torch.Tensor, p2: torch.Tensor, lengths1: Optional[torch.Tensor] = None,
               lengths2: Optional[torch.Tensor] = None, K: int = 1, version: int = -1, return_nn: bool = False,
               return_sorted: bool = True):
    """p1 (B,N,D), p2 (B,M,D) -> KNN(dists (B,N,K), idx (B,N,K) int64, knn (B,N,K,D) or None)."""
    if lengths1 is not None or lengths2 is not None:
        raise NotImplementedError("riggs_b200.knn.knn_points: the reference always passes lengths1 = lengths2 = None")
    if p1.dim() != 3 or p2.dim() != 3 or p1.shape[0] != p2.shape[0] or p1.shape[2] != p2.shape[2]:
        raise ValueError("knn_points: expected p1 (B,N,D) and p2 (B,M,D)")
    if not p1.is_cuda:
        raise _lib.RiggsError("knn_points needs CUDA tensors; riggs_b200 has no CPU path")
    B, N, D = p1.shape
    M = p2.shape[1]
    a, b = f32c(p1.detach()), f32c(p2.detach())
    dists = torch.empty((B, N, K), dtype=torch.float32, device=p1.device)
    idx = torch.empty((B, N, K), dtype=torch.int64, device=p1.device)
    L = _lib.lib()
    with torch.cuda.device(p1.device):
        for i in range(B):
            check(L.rg_knn_points(N, M, D, K, ptr(a[i]), ptr(b[i]), ptr(dists[i]), ptr(idx[i]), stream_ptr()), "rg_knn_points")
    nn = None
    if return_nn or p1.requires_grad or p2.requires_grad:
        nn = torch.gather(p2[:, None].expand(B, N, M, D), 2, idx[..., None].expand(B, N, K, D)) if N * M * D < (1 << 28) \
            else torch.stack([p2[i][idx[i]] for i in range(B)])
        if p1.requires_grad or p2.requires_grad:
            dists = ((p1[:, :, None, :] - nn) ** 2).sum(-1)
    return _KNN(dists, idx, nn if return_nn else None)
