"""Multi-process checker of the library's own all-reduce kernels (run under torchrun by tests/test_gpu_multi.py, or by
hand:  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tests/multi_gpu_check.py).

Every rank fills a symmetric-memory gradient buffer (dist.FlatGrads) with rank-dependent values and sums it over the
ranks three ways: ncclAllReduce on a clone (the comparator), rg_allreduce_p2p (peer loads, rank-ordered adds: must be
bit-identical on every rank and equal to a rank-ordered sum computed with torch) and rg_allreduce_multimem (NVLS: the
switch adds, order unspecified: 1e-6 relative).  rg_allreduce_ranges is checked on both paths with row chunks.
Prints one JSON line per check on rank 0 and exits non-zero on any mismatch."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from riggs_b200.dist import FlatGrads
    ok_all = True
    P = 30_000
    shapes = [("xyz", (P, 3)), ("f_dc", (P, 1, 3)), ("f_rest", (P, 15, 3)), ("opacity", (P, 1)), ("scaling", (P, 3)),
              ("rotation", (P, 4)), ("node_radius", (24,))]
    for mode in ("p2p", "multimem"):
        os.environ["RG_ALLREDUCE"] = mode
        params = [(n, torch.zeros(s, device=dev)) for n, s in shapes]
        flat = FlatGrads(params, device=dev)
        if flat.hdl is None:
            if rank == 0:
                print(json.dumps({"mode": mode, "skipped": "no symmetric memory on this box"}))
            continue
        if mode == "multimem" and not flat._mc_ptr:
            if rank == 0:
                print(json.dumps({"mode": mode, "skipped": "no multicast (NVLS) address on this box"}))
            continue
        g = torch.Generator(device="cpu").manual_seed(100 + rank)
        mine = (torch.randn(flat.total, generator=g) * (1.0 + rank)).to(dev)
        # rank-ordered reference: gather every rank's values, add in rank order
        allv = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allv, mine)
        want = allv[0].clone()
        for r in range(1, world):
            want += allv[r]
        nccl = mine.clone()
        dist.all_reduce(nccl)
        for trial in range(3):  # the kernel runs on a buffer that holds the previous trial's result on every rank
            flat.flat.copy_(mine)
            torch.cuda.synchronize()
            dist.barrier()
            flat.all_reduce()
            torch.cuda.synchronize()
            got = flat.flat.clone()
            if mode == "p2p":
                ok = bool(torch.equal(got, want))
            else:
                ok = float((got - want).abs().max() / want.abs().max()) <= 1e-6
            chk = got.clone()
            dist.broadcast(chk, src=0)
            same = bool(torch.equal(chk, got))  # every rank ends with the same bits (one rank computed each slice)
            err = float((got - nccl).abs().max() / nccl.abs().max())
            t = torch.tensor([int(ok and same and err <= 1e-6)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            ok_all = ok_all and bool(int(t[0]))
            if rank == 0:
                print(json.dumps({"mode": mode, "kernel": "rg_allreduce_" + mode, "world": world, "trial": trial,
                                  "matches_rank_ordered_sum": ok, "identical_on_all_ranks": same,
                                  "max_rel_err_vs_nccl": err, "floats": flat.total}))
        # row-chunked ranges (the overlap path): chunks of rows + the rest with the last chunk
        flat.flat.copy_(mine)
        torch.cuda.synchronize()
        dist.barrier()
        chunks = flat.row_chunks(P, 3)
        for c, (r0, r1) in enumerate(chunks):
            flat.all_reduce_rows(P, r0, r1, with_rest=c == len(chunks) - 1, blocks=16, barrier_after=c == len(chunks) - 1)
        flat.reduced_in_backward = False
        torch.cuda.synchronize()
        got = flat.flat.clone()
        ok = bool(torch.equal(got, want)) if mode == "p2p" else float((got - want).abs().max() / want.abs().max()) <= 1e-6
        t = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok_all = ok_all and bool(int(t[0]))
        if rank == 0:
            print(json.dumps({"mode": mode, "kernel": "rg_allreduce_ranges", "world": world, "chunks": len(chunks), "ok": bool(int(t[0]))}))
        del flat
    ok_all = view_split_check(rank, world, dev) and ok_all
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok_all else 1)


def view_split_check(rank, world, dev) -> bool:
    """One view rendered by all the ranks together (dist.ViewSplit) against the same view rendered by this rank alone:
    image, depth, alpha and radii bit-identical on every rank, every gradient within four times the unsplit step's own run-to-run spread (at least 5e-5; max norm and L2)."""
    from riggs_b200 import skeleton_warp as sw, synth
    from riggs_b200.dist import ViewSplit
    from riggs_b200.fused import WorkspacePool, render_skinned
    from riggs_b200.gaussian_renderer import GaussianParams
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers

    class Pipe:
        debug = False
        convert_SHs_python = False
        compute_cov3D_python = False

    os.environ.pop("RG_ALLREDUCE", None)
    ok_all = True
    for name in ("small", "C2"):
        scene = synth.make_config(name, **({"n_views": 1} if name == "C2" else {}))
        pc = GaussianParams.from_scene(scene, device=dev)
        warp = sw.SkeletonWarp(scene.joints, scene.parents, scene.node_radius_raw, use_skinning_weight_mlp=False,
                               use_template_offsets=False, device=dev)
        cam = scene.cameras[0].to(dev)
        P, W, H = scene.N, cam.image_width, cam.image_height
        bg = torch.tensor([0.1, 0.3, 0.2], device=dev)
        cots = [t.to(dev) for t in synth.seeded_cotangents(H, W)]
        params = [pc._xyz, pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest, warp._node_radius]
        pool = WorkspacePool()
        split = ViewSplit(P, W, H, dev)

        def run(sp):
            pose = {k: v.to(dev).clone().requires_grad_(True) for k, v in scene.poses[0].items()}
            out = render_skinned(cam, pc, warp, pose, Pipe(), bg, pool=pool, split=sp)
            ((out["render"] * cots[0]).sum() + (out["depth"] * cots[1]).sum() + (out["alpha"] * cots[2]).sum()).backward()
            g = [p.grad.clone() for p in params] + [pose["local_rotation"].grad.clone(), pose["global_trans"].grad.clone(),
                                                    out["viewspace_points"].grad.clone()]
            for p in params:
                p.grad = None
            torch.cuda.synchronize()
            return out, g, int(out["status"][0])

        whole, g_whole, n_whole = run(None)
        _, g_again, _ = run(None)  # the unsplit step against itself: float atomics land in another order every run
        spread = max(max(helpers.rel_err(a.cpu(), b.cpu()), helpers.rel_l2(a.cpu(), b.cpu())) for a, b in zip(g_again, g_whole))
        gate = min(max(5e-5, 4 * spread), 1e-3)
        for trial in range(2):
            part, g_part, n_part = run(split)
            same = all(bool(torch.equal(part[k], whole[k])) for k in ("render", "depth", "alpha", "radii"))
            errs = [max(helpers.rel_err(a.cpu(), b.cpu()), helpers.rel_l2(a.cpu(), b.cpu())) for a, b in zip(g_part, g_whole)]
            counts = torch.tensor([n_part], device=dev)
            dist.all_reduce(counts)
            ok = same and max(errs) < gate and int(counts[0]) == n_whole
            t = torch.tensor([int(ok)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            ok_all = ok_all and bool(int(t[0]))
            if rank == 0:
                print(json.dumps({"check": "view_split", "config": name, "world": world, "trial": trial,
                                  "peer_memory": split.images.hdl is not None, "image_bit_identical": same,
                                  "max_grad_err": max(errs), "unsplit_run_to_run": spread, "instances_whole": n_whole, "instances_sum_of_parts": int(counts[0]),
                                  "instances_this_rank": n_part, "ok_all_ranks": bool(int(t[0]))}))
        del split
    return ok_all


if __name__ == "__main__":
    main()
