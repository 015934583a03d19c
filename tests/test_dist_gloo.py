"""CPU-only, world_size 2 over gloo: the host side of the multi-GPU path (riggs_b200/dist.py) -- unit sharding and the
flat gradient buffer that the backward kernels write into and that is all-reduced in place."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers  # noqa: F401
from riggs_b200.dist import FlatGrads, all_reduce_max_, shard_frames, shard_frames_interleaved, shard_views


def test_sharding_covers_every_unit_once():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 7, 8, 256, 257):
            views = sorted(v for r in range(world) for v in shard_views(n, r, world))
            assert views == list(range(n))
            spans = [shard_frames(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
            inter = [shard_frames_interleaved(n, r, world) for r in range(world)]
            assert sorted(f for fr in inter for f in fr) == list(range(n))
            assert max(map(len, inter)) - min(map(len, inter)) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)  # identical replicas on every rank
        P, J = 1000, 24
        named = [("xyz", torch.nn.Parameter(torch.randn(P, 3))), ("features_dc", torch.nn.Parameter(torch.randn(P, 1, 3))),
                 ("features_rest", torch.nn.Parameter(torch.randn(P, 15, 3))), ("opacity", torch.nn.Parameter(torch.randn(P, 1))),
                 ("scaling", torch.nn.Parameter(torch.randn(P, 3))), ("rotation", torch.nn.Parameter(torch.randn(P, 4))),
                 ("node_radius", torch.nn.Parameter(torch.randn(J)))]
        flat = FlatGrads(named)
        assert flat.flat.numel() == P * 59 + J
        # every rank "renders" its views: the kernels would write d(param) into the views; emulate with a per-view value
        views = shard_views(6, rank, world)
        g = torch.Generator().manual_seed(100)
        per_view = [torch.randn(flat.flat.numel(), generator=g) for _ in range(6)]
        for v in views:
            flat.flat += per_view[v]  # views alias the flat buffer
        assert flat.views["xyz"].data_ptr() == flat.flat.data_ptr()
        flat.all_reduce()
        flat.attach()
        expect = sum(per_view)
        ok = torch.allclose(flat.flat, expect, atol=1e-5)
        ok = ok and named[0][1].grad.data_ptr() == flat.views["xyz"].data_ptr()
        ok = ok and torch.allclose(named[6][1].grad, expect[-J:], atol=1e-5)
        # pack fallback (unfused path): grads living elsewhere are copied in
        for _, p in named:
            p.grad = torch.full_like(p, float(rank + 1))
        flat.pack_from_grads()
        flat.all_reduce()
        ok = ok and torch.allclose(flat.flat, torch.full_like(flat.flat, float(sum(range(1, world + 1)))))
        m = all_reduce_max_(torch.tensor([float(rank), 5.0 - rank]))
        ok = ok and m.tolist() == [float(world - 1), 5.0]
        # one view split over the ranks (dist.ViewSplit): rank r owns tile rows y % world == r; the ranks' disjoint image
        # rows and their shares of the per-Gaussian accumulators are summed in place on every rank
        from riggs_b200.dist import ViewSplit
        sp = ViewSplit(P, 64, 48, "cpu")
        ok = ok and (sp.index, sp.count, sp.emulated) == (rank, world, False)
        color, depth, alpha = sp.image_views()
        ok = ok and color.shape == (3, 48, 64) and depth.shape == (1, 48, 64) and alpha.data_ptr() == sp.images.buf[4 * 48 * 64:].data_ptr()
        rows = (torch.arange(48) // 16) % world == rank
        sp.images.buf.zero_()
        color[:, rows] = 1.0 + rank
        sp.sum_images()
        want = (1.0 + (torch.arange(48) // 16) % world).float()[None, :, None].expand(3, 48, 64)
        ok = ok and torch.equal(color, want)
        acc = torch.full((10 * P,), float(rank + 1))
        sp.sum_grad2d(acc)
        ok = ok and torch.equal(acc, torch.full((10 * P,), float(sum(range(1, world + 1)))))
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_flat_gradient_all_reduce_world2_gloo():
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}
