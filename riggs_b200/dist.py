"""Multi-GPU plumbing of the hot path: one process per GPU (torchrun), torch.distributed for the collective.

The path shards by view / frame: a view's skin + render forward/backward touches no other
view's data.  Training steps are data-parallel over views -- every rank holds a full replica of the Gaussian and
skeleton parameters, renders its views, and the parameter gradients are summed by ONE all-reduce over a flat fp32
buffer that the backward kernels write into directly (no packing copy).  Novel-pose sequences are sharded by
contiguous frame ranges with no collective at all.  The reference has no distributed code; this module is new.
"""
from __future__ import annotations

import os
from typing import Dict, List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Views of a training batch handled by `rank`: r, r + world, r + 2*world, ..."""
    return list(range(rank, n_views, world))


def shard_frames(n_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [begin, end) frame range of `rank` for forward-only sequences; sizes differ by at most one."""
    base, extra = divmod(n_frames, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_frames_interleaved(n_frames: int, rank: int, world: int) -> List[int]:
    """Frames rank, rank + world, rank + 2*world, ...: the cost of a frame (its instance count) drifts slowly along a
    pose sequence, so contiguous ranges leave the ranks unevenly loaded (measured on C5, 2 GPUs: 1.08 vs 1.63 ms per
    frame) while interleaved frames give every rank the same mix.  Frames keep their index, so the output order is
    unaffected."""
    return list(range(rank, n_frames, world))


class FlatGrads:
    """One flat fp32 gradient buffer with a named view per parameter.

    Layout = the parameters in the order given (for the Gaussians: xyz 3 | f_dc 3 | f_rest 45 | opacity 1 | scaling 3 |
    rotation 4 per Gaussian-major tensor, then the skeleton parameters).  The fused backward writes its outputs straight
    into the views (``render_skinned(..., grad_out=flat.views)``), ``all_reduce`` sums the buffer over the ranks in
    place, ``attach`` makes every ``param.grad`` a view of the reduced buffer."""

    def __init__(self, named_params: Sequence[Tuple[str, torch.Tensor]], device=None, peer_memory: bool = True):
        """peer_memory: with NCCL on 2, 4 or 8 GPUs of one node the buffer is allocated as symmetric memory (every
        rank's copy mapped into every process over NVLink) and ``all_reduce`` runs this library's own kernel
        (rg_allreduce_p2p) between two device-side barriers instead of ncclAllReduce; anything else uses NCCL / gloo."""
        self.names = [n for n, _ in named_params]
        self.params = [p for _, p in named_params]
        device = torch.device(device if device is not None else self.params[0].device)
        total = sum(p.numel() for p in self.params)
        self.total = total
        self.hdl = None
        self._peer_ptrs = None
        self._mc_ptr = 0
        self.reduced_in_backward = False  # set by all_reduce_rows: the buffer already holds the sum
        # Opt-in: sum the gradient chunk by chunk on a second stream while the backward computes later chunks
        # (fused.render_skinned(reduce=...)).  Measured on 2 x B200 (C0): 0.725-0.731 ms/step against 0.72 ms with the
        # single all-reduce after the backward -- the issue-bound per-Gaussian backward and the link-bound sum slow each
        # other down by as much as the overlap hides, so it stays off.
        self.overlap = os.environ.get("RG_DP_OVERLAP", "0") == "1"
        self.overlap_blocks = int(os.environ.get("RG_DP_BLOCKS", "32"))   # CTAs of the reduce kernel running beside the backward
        self.overlap_chunks = int(os.environ.get("RG_DP_CHUNKS", "4"))
        padded = (total + 3) // 4 * 4  # the peer kernel moves 16-byte words
        if os.environ.get("RG_ALLREDUCE", "") == "nccl":
            peer_memory = False
        if peer_memory and device.type == "cuda":
            self._try_symmetric(padded, device)
        if self.hdl is None:
            self._storage = torch.zeros(padded, dtype=torch.float32, device=device)
        self.flat = self._storage[:total]
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for n, p in named_params:
            k = p.numel()
            self.views[n] = self.flat[off:off + k].view(p.shape)
            off += k

    def _try_symmetric(self, padded: int, device):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_backend() != "nccl":
            return
        if dist.get_world_size() not in (2, 4, 8):
            return
        try:
            import torch.distributed._symmetric_memory as symm_mem
            group = dist.group.WORLD.group_name
            t = symm_mem.empty(padded, dtype=torch.float32, device=device)
            hdl = symm_mem.rendezvous(t, group)
            t.zero_()
            import ctypes as C
            off = t.data_ptr() - hdl.buffer_ptrs[hdl.rank]
            self._peer_ptrs = (C.c_void_p * hdl.world_size)(*[int(hdl.buffer_ptrs[r]) + off for r in range(hdl.world_size)])
            # NVLS multicast address of the same buffer: the switch reduces (multimem.ld_reduce) and replicates
            # (multimem.st), which moves (N+1)/N of the buffer per link direction instead of 2(N-1)/N -- measured on
            # B200: slower than the peer kernel at N=2 (0.19 vs 0.11 ms), faster at N=8 (0.15 vs 0.22 ms)
            want_mc = {"multimem": True, "p2p": False}.get(os.environ.get("RG_ALLREDUCE", ""), hdl.world_size >= 8)
            mc = int(hdl.multicast_ptr) if want_mc else 0
            self._mc_ptr = mc + off if mc else 0
            self._storage, self.hdl = t, hdl
        except Exception as e:  # no peer access / no symmetric-memory support: NCCL does the job
            import warnings
            warnings.warn(f"riggs_b200: symmetric memory unavailable ({e}); gradients are all-reduced by NCCL")
            self.hdl = None

    def zero_(self):
        self.flat.zero_()

    def pack_from_grads(self):
        """Fallback for paths whose backward does not write into the views: copy param.grad into the buffer."""
        for n, p in zip(self.names, self.params):
            if p.grad is None:
                self.views[n].zero_()
            elif p.grad.data_ptr() != self.views[n].data_ptr():
                self.views[n].copy_(p.grad)

    def all_reduce(self, average: bool = False, async_op: bool = False):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return None
        if self.reduced_in_backward:  # summed chunk by chunk while the backward ran
            self.reduced_in_backward = False
            if average:
                self.flat.div_(dist.get_world_size())
            return None
        if self.hdl is not None and not async_op:
            from . import _lib
            hdl = self.hdl
            hdl.barrier(channel=0)  # every rank's backward has written its copy
            if self._mc_ptr:
                _lib.check(_lib.lib().rg_allreduce_multimem(self._mc_ptr, self._storage.numel(), hdl.rank, hdl.world_size,
                                                            _lib.stream_ptr()), "rg_allreduce_multimem")
            else:
                _lib.check(_lib.lib().rg_allreduce_p2p(self._peer_ptrs, self._storage.numel(), hdl.rank, hdl.world_size,
                                                       _lib.stream_ptr()), "rg_allreduce_p2p")
            hdl.barrier(channel=1)  # every rank's slice has arrived in every copy
            if average:
                self.flat.div_(dist.get_world_size())
            return None
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=async_op)
        if average and not async_op:
            self.flat.div_(dist.get_world_size())
        return work

    # ---- all-reduce overlapped with the backward that fills the buffer ------------------------------------
    def can_overlap(self, n_rows: int) -> bool:
        """True when the buffer is peer-mapped and every per-row parameter (first dimension n_rows) and every other
        parameter starts and ends on a 16-byte word, so that chunks of rows can be summed while later rows are still
        being computed (``fused.render_skinned(..., reduce=flat)``)."""
        if self.hdl is None or n_rows % 4:
            return False
        off = 0
        n_ranges = 0
        for p in self.params:
            if off % 4 or p.numel() % 4:
                return False
            off += p.numel()
            n_ranges += 1
        return n_ranges <= 8

    def row_chunks(self, n_rows: int, n_chunks: int = 0):
        """Row ranges of about equal size whose starts are multiples of 256 (the backward's CTA granularity)."""
        n_chunks = n_chunks or self.overlap_chunks
        step = max(256, (n_rows + n_chunks - 1) // n_chunks // 256 * 256)
        out, r = [], 0
        while r < n_rows:
            e = min(n_rows, r + step)
            if n_rows - e < 256:
                e = n_rows
            out.append((r, e))
            r = e
        return out

    def all_reduce_rows(self, n_rows: int, r0: int, r1: int, with_rest: bool, blocks: int = 0, barrier_after: bool = False):
        """Device-side barrier, then the sum of rows [r0, r1) of every per-row parameter (plus, with_rest, of the other
        parameters) on the current stream.  The caller orders this after the kernels that wrote those rows; successive
        calls need no barrier in between (the next call's leading barrier also fences this call's kernel), the last one
        passes barrier_after=True."""
        import ctypes as C
        from . import _lib
        blocks = blocks or self.overlap_blocks
        offs, lens, off = [], [], 0
        for p in self.params:
            k = p.numel()
            if p.dim() > 0 and p.shape[0] == n_rows:
                w = k // n_rows
                offs.append(off + r0 * w)
                lens.append((r1 - r0) * w)
            elif with_rest:
                offs.append(off)
                lens.append(k)
            off += k
        hdl = self.hdl
        hdl.barrier(channel=0)
        if offs:
            n = len(offs)
            _lib.check(_lib.lib().rg_allreduce_ranges(self._peer_ptrs, self._mc_ptr or None, n, (C.c_size_t * n)(*offs),
                                                      (C.c_size_t * n)(*lens), hdl.rank, hdl.world_size, blocks,
                                                      _lib.stream_ptr()), "rg_allreduce_ranges")
        if barrier_after:
            hdl.barrier(channel=0)
        self.reduced_in_backward = True

    def attach(self):
        for n, p in zip(self.names, self.params):
            p.grad = self.views[n]


class PeerSum:
    """A fp32 buffer of `numel` values that ``sum_()`` adds up over the ranks in place: symmetric memory + this library's
    own all-reduce kernels (the same two as ``FlatGrads``) on 2 / 4 / 8 GPUs of a node under NCCL, ``dist.all_reduce``
    otherwise (gloo on the CPU for the tests), nothing at world size 1."""

    def __init__(self, numel: int, device, peer_memory: bool = True):
        self.numel = int(numel)
        padded = (self.numel + 3) // 4 * 4
        self.hdl = None
        self._peer_ptrs = None
        self._mc_ptr = 0
        device = torch.device(device)
        if (peer_memory and device.type == "cuda" and dist.is_available() and dist.is_initialized()
                and dist.get_backend() == "nccl" and dist.get_world_size() in (2, 4, 8)
                and os.environ.get("RG_ALLREDUCE", "") != "nccl"):
            try:
                import ctypes as C
                import torch.distributed._symmetric_memory as symm_mem
                t = symm_mem.empty(padded, dtype=torch.float32, device=device)
                hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
                t.zero_()
                off = t.data_ptr() - hdl.buffer_ptrs[hdl.rank]
                self._peer_ptrs = (C.c_void_p * hdl.world_size)(*[int(hdl.buffer_ptrs[r]) + off for r in range(hdl.world_size)])
                want_mc = {"multimem": True, "p2p": False}.get(os.environ.get("RG_ALLREDUCE", ""), hdl.world_size >= 8)
                mc = int(hdl.multicast_ptr) if want_mc else 0
                self._mc_ptr = mc + off if mc else 0
                self._storage, self.hdl = t, hdl
            except Exception as e:
                import warnings
                warnings.warn(f"riggs_b200: symmetric memory unavailable ({e}); using the process group's all-reduce")
                self.hdl = None
        if self.hdl is None:
            self._storage = torch.zeros(padded, dtype=torch.float32, device=device)
        self.buf = self._storage[:self.numel]

    def sum_(self):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return
        if self.hdl is None:
            dist.all_reduce(self.buf, op=dist.ReduceOp.SUM)
            return
        from . import _lib
        hdl = self.hdl
        hdl.barrier(channel=0)
        if self._mc_ptr:
            _lib.check(_lib.lib().rg_allreduce_multimem(self._mc_ptr, self._storage.numel(), hdl.rank, hdl.world_size,
                                                        _lib.stream_ptr()), "rg_allreduce_multimem")
        else:
            _lib.check(_lib.lib().rg_allreduce_p2p(self._peer_ptrs, self._storage.numel(), hdl.rank, hdl.world_size,
                                                   _lib.stream_ptr()), "rg_allreduce_p2p")
        hdl.barrier(channel=1)


class ViewSplit:
    """ONE view rendered by all the ranks together (last bullet): what cuts the latency of a single
    frame, where sharding by view only raises throughput.  The reference has nothing like it.

    Tile row y of the 16-pixel tile grid belongs to rank y % count (interleaved: the instances of a frame sit in the
    middle rows, contiguous bands would leave the outer ranks idle).  Every rank skins, projects and depth-sorts all the
    Gaussians (that part is per-Gaussian streaming and cheap next to an exchange of the 64-byte splat records), then bins
    and blends only its rows.  Forward: the ranks' disjoint rows of (colour, depth, alpha) are summed into the full image
    on every rank.  Backward: every rank runs the blend backward over its rows, the per-Gaussian screen-space
    accumulators (10 floats per Gaussian) are summed over the ranks, and the per-Gaussian + skeleton backward then runs
    replicated, so every rank ends with the full, identical parameter gradients: no further all-reduce.

    Every forward and backward contains one collective, so the ranks must issue the same sequence of frames: size the
    instance buffers on every rank before capturing graphs (``HotPath.measure_instances`` / ``fused.measure_instances``) --
    a re-capture that an overflow triggers on one rank alone would leave the others waiting in a barrier.

    ``index`` / ``count`` default to the process group's rank / world size; explicit values without a process group give
    one part of an emulated split on a single GPU (tests sum the parts themselves through ``images`` / ``grad2d``)."""

    def __init__(self, P: int, width: int, height: int, device, index=None, count=None, peer_memory: bool = True):
        live = dist.is_available() and dist.is_initialized()
        self.count = int(count if count is not None else (dist.get_world_size() if live else 1))
        self.index = int(index if index is not None else (dist.get_rank() if live else 0))
        self.P, self.width, self.height = int(P), int(width), int(height)
        self.images = PeerSum(5 * self.height * self.width, device, peer_memory)  # colour (3), depth, alpha
        self.grad2d = PeerSum(10 * self.P, device, peer_memory)
        self.emulated = not live or (count is not None and count != dist.get_world_size())
        self.exchange = None  # emulated split only: callable(split, acc) standing in for the sum over the ranks (tests)

    def image_views(self):
        H, W = self.height, self.width
        b = self.images.buf
        return b[:3 * H * W].view(3, H, W), b[3 * H * W:4 * H * W].view(1, H, W), b[4 * H * W:].view(1, H, W)

    def sum_images(self):
        if not self.emulated:
            self.images.sum_()

    def sum_grad2d(self, acc: torch.Tensor):
        """acc: the workspace's accumulator array (10 P floats), summed over the ranks in place."""
        if self.emulated:
            if self.exchange is not None:
                self.exchange(self, acc)
            return
        self.grad2d.buf.copy_(acc)
        self.grad2d.sum_()
        acc.copy_(self.grad2d.buf)


def all_reduce_max_(t: torch.Tensor) -> torch.Tensor:
    """Rank-wise maximum (e.g. max_radii2D of the densification statistics, scene/gaussian_model.py:516-518)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t
