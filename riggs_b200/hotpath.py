"""One iteration of the hot path as the reference's trainer runs it (train_rig.py:535-554 minus SSIM and the
regularisers): pose -> forward kinematics -> skin all Gaussians -> render one view -> L1 loss -> backward to
every Gaussian parameter and to the skeleton inputs.  Used by bench.py, the smoke test and the multi-GPU path.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import skeleton_warp as sw
from .gaussian_renderer import GaussianParams, render
from .loss_utils import image_loss_and_grad, l1_loss, l1_loss_and_grad  # noqa: F401


class _Pipe:
    debug = False
    convert_SHs_python = False
    compute_cov3D_python = False


class HotPath:
    def __init__(self, scene, device="cuda", use_mlp_weight: bool = False, use_offsets: bool = False,
                 fused: bool = True, lambda_dssim: float = 0.0, mlps: bool = False, inference: bool = False):
        """fused=True: one C-ABI call per direction (rg_skinned_forward/backward, no host sync);
        fused=False: the reference's call sequence (deform_by_pose, then render()) on the same kernels.
        mlps=True: the reference's shipped configuration (scripts/run_demo.py:32, --use_skinning_weight_mlp
        --use_template_offsets): sigmoid(WeightMLP(x)) modulates the skinning weights and detail_net(x, pose) adds a
        per-Gaussian offset, both evaluated and differentiated every step (skeleton_warp.py:56-61, 98-104, 152-153).
        inference=True: no gradients are wanted (novel-pose rendering): the pose-independent WeightMLP output is
        evaluated once and reused for every frame."""
        self.fused = fused
        self.lambda_dssim = float(lambda_dssim)  # > 0: the trainer's L1 + SSIM image loss (train_rig.py:508-509)
        self.device = torch.device(device)
        self.pc = GaussianParams.from_scene(scene, device=device)
        self.warp = sw.SkeletonWarp(scene.joints, scene.parents, scene.node_radius_raw,
                                    use_skinning_weight_mlp=False, use_template_offsets=False, device=device)
        self.mlp_weight = scene.mlp_weight.to(device) if (use_mlp_weight and scene.mlp_weight is not None) else None
        self.offsets = scene.template_offsets.to(device) if (use_offsets and scene.template_offsets is not None) else None
        self.mlps = bool(mlps)
        self.inference = bool(inference)
        self.weight_mlp = self.detail_net = None
        self.fused_weight_mlp = self.fused_detail_net = None
        self._weight_cache = None
        if self.mlps:
            torch.manual_seed(4321)  # the CPU arm of bench.py draws the same weights
            J = scene.joints.shape[0]
            wm, dn = sw.WeightMLP(3, J - 1), sw.DeformMLP(4 * J)
            with torch.no_grad():
                dn.gaussian_warp.weight.normal_(0, 1e-3)  # the reference's 1e-5 start would leave the offsets at ~0
            self.weight_mlp, self.detail_net = wm.to(device), dn.to(device)
            if mlps != "torch":  # the tensor-core kernels (csrc/mlp.cu); "torch": the eager fp32 modules, for comparison
                from .mlp import FusedSkipMLP
                self.fused_weight_mlp, self.fused_detail_net = FusedSkipMLP(self.weight_mlp), FusedSkipMLP(self.detail_net)
        self.pipe = _Pipe()
        self.bg = torch.zeros(3, device=device)
        self.isotropic = self.pc.use_isotropic_gs
        self.flat = None
        self.pool = None  # fused.WorkspacePool override (GraphedStep keeps a private one)
        self.split = None  # dist.ViewSplit: every view is rendered by all the ranks together (enable_view_split)

    def measure_instances(self, cams, poses, pool) -> int:
        """Renders every (camera, pose) pair once without gradients and sizes `pool`'s instance buffers for the largest
        instance count seen (fused.WorkspacePool.reserve); returns that count."""
        old, self.pool = self.pool, pool
        r_max = 0
        P, dev = self.pc._xyz.shape[0], self.pc._xyz.device
        try:
            with torch.no_grad():
                for cam, pose in zip(cams, poses):
                    out = self.forward(cam, pose)
                    torch.cuda.current_stream().synchronize()
                    r = int(out["status"][0])
                    if int(out["status"][1]) != 0:
                        pool.reserve(P, int(cam.image_width), int(cam.image_height), dev, r)
                    r_max = max(r_max, r)
        finally:
            self.pool = old
        if cams:
            pool.reserve(P, int(cams[0].image_width), int(cams[0].image_height), dev, r_max)
        return r_max

    def mlp_parameters(self):
        if not self.mlps:
            return []
        return list(self.weight_mlp.parameters()) + list(self.detail_net.parameters())

    def _front_end(self, pose):
        """The per-Gaussian MLP outputs of this step: (mlp_weight (P,J-1) or None, offsets (P,3) or None)."""
        if not self.mlps:
            return self.mlp_weight, self.offsets
        x = self.pc._xyz.detach()
        if self.fused_weight_mlp is not None:
            q = pose["local_rotation"].detach().reshape(-1)
            wmlp, dnet = self.fused_weight_mlp, (lambda xx, qq: self.fused_detail_net(xx, qq))
        else:
            q = pose["local_rotation"].detach().reshape(-1)[None].expand(x.shape[0], -1)
            wmlp, dnet = self.weight_mlp, self.detail_net
        if self.inference:
            with torch.no_grad():
                if self._weight_cache is None:
                    self._weight_cache = wmlp(x)
                return self._weight_cache, dnet(x, q)
        return wmlp(x), dnet(x, q)

    def named_parameters(self):
        pc = self.pc
        return [("xyz", pc._xyz), ("features_dc", pc._features_dc), ("features_rest", pc._features_rest),
                ("opacity", pc._opacity), ("scaling", pc._scaling), ("rotation", pc._rotation),
                ("node_radius", self.warp._node_radius)]

    def enable_flat_grads(self):
        """Data-parallel mode: the fused backward writes the parameter gradients straight into one flat buffer
        (dist.FlatGrads) that is all-reduced in place; ``reduce_grads()`` then exposes it through ``param.grad``."""
        from .dist import FlatGrads
        self.flat = FlatGrads(self.named_parameters(), device=self.device)
        return self.flat

    def enable_view_split(self, width: int, height: int, index=None, count=None):
        """Latency mode: all the ranks render the SAME view, each binning and blending every count-th tile row
        (dist.ViewSplit).  Images and gradients come out complete and identical on every rank, so there is no flat
        buffer and no gradient all-reduce; the two sums that replace it (image rows, 10 floats per Gaussian) are part of
        the forward / backward and of any graph captured around them."""
        from .dist import ViewSplit
        if not self.fused:
            raise ValueError("the view split lives in the fused entry points: HotPath(fused=True)")
        self.split = ViewSplit(self.pc._xyz.shape[0], width, height, self.device, index=index, count=count)
        if self.split.count <= 1:
            self.split = None
        return self.split

    def reduce_grads(self):
        if self.flat is None:
            return
        if not self.fused:
            self.flat.pack_from_grads()
        self.flat.all_reduce()
        self.flat.attach()

    def parameters(self):
        pc = self.pc
        return [pc._xyz, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation,
                self.warp._node_radius]

    def zero_grad(self):
        for p in self.parameters():
            p.grad = None
        for p in self.mlp_parameters():
            p.grad = None

    def forward(self, cam, pose: Dict[str, torch.Tensor], accumulate: bool = False, last_view: bool = True):
        """pose: {'local_rotation': (J,4), 'global_trans': (3,)} on the device (may require grad).
        accumulate: this view's backward ADDS to the flat gradient buffer (a later view of a multi-view batch)."""
        mlp_weight, offsets = self._front_end(pose)
        if self.fused:
            from .fused import render_skinned
            return render_skinned(cam, self.pc, self.warp, pose, self.pipe, self.bg, mlp_weight=mlp_weight,
                                  offsets=offsets, grad_out=self.flat.views if self.flat is not None else None,
                                  pool=self.pool, accumulate=accumulate, split=self.split,
                                  reduce=self.flat if (getattr(self.flat, "overlap", False) and last_view) else None)
        if accumulate:
            raise ValueError("multi-view accumulation needs HotPath(fused=True)")
        joints = self.warp.nodes[:, :3]
        G, node_rot, d_nodes = sw.fk(pose["local_rotation"], pose["global_trans"], joints, self.warp.parents_i32)
        mask = self.pc._motion_mask.reshape(-1) if self.pc._motion_mask is not None else None
        d_xyz, d_rot, _ = sw.skin(self.pc.get_xyz, joints, self.warp.parents_i32, self.warp._node_radius, G, node_rot,
                                  pose["global_trans"], mask, mlp_weight, offsets, 0, False)
        if self.isotropic:  # the trainer zeroes d_rotation for isotropic Gaussians (train_rig.py:422-423)
            d_rot = 0.0
        return render(cam, self.pc, self.pipe, self.bg, d_xyz, d_rot, 0.0)

    def step_views(self, cams, poses, targets) -> torch.Tensor:
        """One optimiser step's worth of views on this rank (a multi-view batch: BASELINE configs with 4 or 8 views per
        step): the first view's backward overwrites the flat gradient buffer, every later view's adds to it, so no
        zero-fill pass and no per-view gradient tensors exist.  Returns the per-view losses (V,)."""
        if len(cams) > 1 and self.flat is None:
            self.enable_flat_grads()
        n = len(cams)
        losses = [self.step(c, p, t, accumulate=i > 0, last_view=i == n - 1) for i, (c, p, t) in enumerate(zip(cams, poses, targets))]
        return torch.stack([l.reshape(()) for l in losses])

    def step(self, cam, pose, target: torch.Tensor, accumulate: bool = False, last_view: bool = True) -> torch.Tensor:
        """Forward + backward of one view; returns the (device) loss.  Gradients accumulate in .grad."""
        out = self.forward(cam, pose, accumulate=accumulate, last_view=last_view)
        # radii / viewspace_points / status of the last view (densification statistics, overflow check); nothing that
        # carries a grad_fn is kept: a graph held across steps keeps its AccumulateGrad nodes (and their stream) alive
        self.last_out = {k: out[k] for k in ("radii", "viewspace_points", "status")}
        if self.lambda_dssim > 0.0:
            loss, grad, _l1, _ssim = image_loss_and_grad(out["render"], target, self.lambda_dssim)
        else:
            loss, grad = l1_loss_and_grad(out["render"], target)  # one kernel: the loss and d loss / d image
        out["render"].backward(grad)
        return loss


class GraphedStep:
    """``HotPath.step`` for a fixed (P, J, width, height) captured once as a CUDA graph and replayed per view.

    A step is ~30 short launches (kernels of this library, the CUB sort passes, the loss' element-wise kernels);
    replayed from a graph they run back to back without the per-launch gaps of a Python-driven stream.
    Inputs are copied into static device buffers (``cam`` matrices, pose, target image); gradients are OVERWRITTEN
    by every replay (param.grad / ``pose_grads``), not accumulated.  The instance-list capacity check of the fused
    path runs one replay late, exactly as in eager mode; an overflow re-captures with a larger buffer.
    """

    def __init__(self, hp: HotPath, cam, pose, target, warmup: int = 3, pool=None):
        """cam / pose / target: one view, or lists of V views = the views this rank renders per optimiser step (their
        parameter gradients are summed inside the graph: first view overwrites, later views add)."""
        if not hp.fused:
            raise ValueError("GraphedStep needs HotPath(fused=True): the unfused path synchronises with the host")
        from .fused import WorkspacePool
        import copy
        self.hp = hp
        dev = hp.device
        self.pool = pool if pool is not None else WorkspacePool()  # a shared pool = one workspace for several graphs
        multi = isinstance(cam, (list, tuple))
        cams, poses, targets = (list(cam), list(pose), list(target)) if multi else ([cam], [pose], [target])
        if len(cams) > 1 and hp.flat is None:
            hp.enable_flat_grads()
        self.cams, self.poses, self.targets = [], [], []
        for c, p, t in zip(cams, poses, targets):
            cc = copy.copy(c)
            cc.world_view_transform = c.world_view_transform.to(dev).clone()
            cc.full_proj_transform = c.full_proj_transform.to(dev).clone()
            cc.camera_center = c.camera_center.to(dev).clone()
            self.cams.append(cc)
            self.poses.append({k: v.detach().to(dev).clone().requires_grad_(True) for k, v in p.items()})
            self.targets.append(t.to(dev).clone())
        self.cam, self.pose, self.target = self.cams[0], self.poses[0], self.targets[0]  # the single-view names
        self.n_views = len(self.cams)
        self.loss: Optional[torch.Tensor] = None
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._event: Optional[torch.cuda.Event] = None
        self._warmup = max(int(warmup), 1)
        self._capture()

    def _one(self):
        hp = self.hp
        hp.zero_grad()
        for pose in self.poses:
            for v in pose.values():
                v.grad = None
        old, hp.pool = getattr(hp, "pool", None), self.pool
        try:
            if self.n_views == 1:
                loss = hp.step(self.cam, self.pose, self.target)
            else:
                loss = hp.step_views(self.cams, self.poses, self.targets)
            if self._reduce_in_graph():
                # data parallel over a peer-mapped gradient buffer: the cross-rank sum (two device-side barriers and this
                # library's all-reduce kernel) is part of the captured step, not an eager call after each replay
                hp.flat.all_reduce()
            return loss
        finally:
            hp.pool = old

    def _reduce_in_graph(self) -> bool:
        flat = self.hp.flat
        return flat is not None and getattr(flat, "hdl", None) is not None

    def _capture(self):
        from . import _lib
        s = torch.cuda.Stream(device=self.hp.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):  # sizes the workspace (first frame is synchronous) and warms the allocator
            for _ in range(self._warmup):
                self._one()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.lib().rg_launch_count()
        with torch.cuda.graph(self.graph):
            self.loss = self._one()
        self.launches_per_replay = int(_lib.lib().rg_launch_count() - n0)
        self.grads = {n: p.grad for n, p in self.hp.named_parameters()}  # static: rewritten by every replay
        # data parallel with a peer-mapped gradient buffer: the all-reduce was captured with the backward
        self._reduces_in_graph = bool(self.hp.flat is not None and (self.hp.flat.reduced_in_backward or self._reduce_in_graph()))
        self._ws = self.pool.free and next(iter(self.pool.free.values()))[0]
        self._captured_binning = self._ws.binning.data_ptr() if self._ws else None
        self._event = None

    @property
    def pose_grads(self) -> Dict[str, torch.Tensor]:
        return {k: v.grad for k, v in self.pose.items()}

    @property
    def overflowed(self) -> bool:
        """True when the last replay produced more instances than the instance buffer holds (its image and gradients
        are incomplete and must not reach the optimiser).  Synchronises with the replay; the next ``step`` re-captures
        with a larger buffer."""
        if self._event is None or self._ws is None:
            return False
        self._event.synchronize()
        return int(self._ws.status[1]) != 0

    def set_inputs(self, cam=None, pose=None, target=None, stream=None) -> None:
        """Copies new inputs (one view or a list of V views, device or pinned host tensors) into the static buffers."""
        def as_list(x):
            return list(x) if isinstance(x, (list, tuple)) else [x]
        with torch.no_grad():
            if cam is not None:
                for dst, src in zip(self.cams, as_list(cam)):
                    dst.world_view_transform.copy_(src.world_view_transform, non_blocking=True)
                    dst.full_proj_transform.copy_(src.full_proj_transform, non_blocking=True)
                    dst.camera_center.copy_(src.camera_center, non_blocking=True)
            if pose is not None:
                for dst, src in zip(self.poses, as_list(pose)):
                    for k, v in src.items():
                        dst[k].copy_(v, non_blocking=True)
            if target is not None:
                for dst, src in zip(self.targets, as_list(target)):
                    dst.copy_(src, non_blocking=True)

    def prefetch(self, cam=None, pose=None, target=None) -> None:
        """Starts copying the NEXT step's inputs (typically pinned host tensors) into staging buffers on a side
        stream, so that the transfer overlaps the step that is running; the next ``step()`` without arguments picks
        them up (a device-to-device copy of a few microseconds) before it replays."""
        dev = self.hp.device
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._stage_cam = [torch.empty_like(t) for t in (self.cam.world_view_transform, self.cam.full_proj_transform,
                                                             self.cam.camera_center)]
            self._stage_pose = {k: torch.empty_like(v) for k, v in self.pose.items()}
            self._stage_target = torch.empty_like(self.target)
            self._consumed = None
        cs = self._copy_stream
        if self._consumed is not None:
            cs.wait_event(self._consumed)  # the previous staged set has been moved into the static inputs
        self._staged = (cam is not None, pose is not None, target is not None)
        with torch.cuda.stream(cs), torch.no_grad():
            if cam is not None:
                for dst, src in zip(self._stage_cam, (cam.world_view_transform, cam.full_proj_transform, cam.camera_center)):
                    dst.copy_(src, non_blocking=True)
            if pose is not None:
                for k, v in pose.items():
                    self._stage_pose[k].copy_(v, non_blocking=True)
            if target is not None:
                self._stage_target.copy_(target, non_blocking=True)
            self._ready = torch.cuda.Event()
            self._ready.record(cs)

    def _take_staged(self):
        ready = getattr(self, "_ready", None)
        if ready is None:
            return
        torch.cuda.current_stream().wait_event(ready)
        has_cam, has_pose, has_target = self._staged
        with torch.no_grad():
            if has_cam:
                for dst, src in zip((self.cam.world_view_transform, self.cam.full_proj_transform, self.cam.camera_center),
                                    self._stage_cam):
                    dst.copy_(src, non_blocking=True)
            if has_pose:
                for k, v in self._stage_pose.items():
                    self.pose[k].copy_(v, non_blocking=True)
            if has_target:
                self.target.copy_(self._stage_target, non_blocking=True)
        self._consumed = torch.cuda.Event()
        self._consumed.record()
        self._ready = None

    def step(self, cam=None, pose=None, target=None) -> torch.Tensor:
        """Copies whatever is given (or was staged by ``prefetch``) into the static inputs, replays the graph and
        returns the (device) loss."""
        ws = self._ws
        if ws is not None and self._event is not None and self._event.query() and int(ws.status[1]) != 0:
            import warnings
            warnings.warn("riggs_b200: the previous frame overflowed the instance buffer; re-capturing with a larger one")
            from . import _lib
            ws.grow_binning(int(_lib.lib().rg_binning_bytes(int(int(ws.status[0]) * self.pool.headroom) + 256)))
            self.pool.note_capacity(ws, self.hp.device)
            self._capture()
        elif ws is not None and ws.binning.data_ptr() != self._captured_binning:
            self._capture()  # another graph on the same (shared) workspace grew the instance buffer
        self._take_staged()
        self.set_inputs(cam, pose, target)
        self.graph.replay()
        if self.hp.flat is not None:
            self.hp.flat.attach()  # param.grad = views of the flat buffer the replay wrote
        if self._reduces_in_graph:
            self.hp.flat.reduced_in_backward = True
        for n, p in self.hp.named_parameters():
            if self.grads[n] is not None and p.grad is not self.grads[n]:
                p.grad = self.grads[n]
        self._event = torch.cuda.Event()
        self._event.record()
        return self.loss


class PipelinedSteps:
    """Input / output pipeline around ``GraphedStep`` for callers whose per-step inputs live in HOST memory
    (a data loader's pinned target image, camera and pose) and who read every step's loss on the host.

    Two captured graphs (slots) with their own static input buffers alternate: while slot ``i % 2``
    computes step i, ``feed`` copies step i+1's host inputs straight into the other slot's static buffers on a
    copy stream (no staging copy on the compute stream), and step i's loss travels to a pinned host word that
    ``result`` reads one step late.  The compute stream therefore always has the next replay queued; nothing
    waits for the host.  Gradients are those of the slot replayed last (``param.grad``), as with ``GraphedStep``.
    """

    def __init__(self, hp: HotPath, cam, pose, target, warmup: int = 3, slots: int = 2, share_workspace: bool = True,
                 pool=None):
        from .fused import WorkspacePool
        self.hp = hp
        # the slots run one after the other on one stream, so they can work in the same geometry / binning / image
        # buffers (and keep that state in L2 from step to step); only the inputs are per slot
        pool = pool if pool is not None else (WorkspacePool() if share_workspace else None)
        self.slots = [GraphedStep(hp, cam, pose, target, warmup=warmup, pool=pool) for _ in range(slots)]
        self.launches_per_replay = self.slots[0].launches_per_replay
        self._copy_stream = torch.cuda.Stream(device=hp.device)
        self._out_stream = torch.cuda.Stream(device=hp.device)
        self._ready = [None] * slots   # inputs of the slot's next step have arrived
        self._done = [None] * slots    # the slot's last replay (and its loss copy) has finished
        self._loss_host = [torch.zeros(self.slots[0].n_views).pin_memory() for _ in range(slots)]
        self.overflowed_steps = []  # steps whose frame overflowed the instance buffer (see result())
        self._fed = 0
        self._run = 0

    def feed(self, cam=None, pose=None, target=None) -> None:
        """Queues the host -> device copies of the NEXT unfed step's inputs (pinned tensors make them asynchronous)."""
        k = self._fed % len(self.slots)
        g, cs = self.slots[k], self._copy_stream
        if self._done[k] is not None:
            cs.wait_event(self._done[k])  # the slot's previous step no longer reads these buffers
        with torch.cuda.stream(cs):
            g.set_inputs(cam, pose, target)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._ready[k] = ev
        self._fed += 1

    def launch(self) -> int:
        """Replays the next step on the current stream (after its inputs have arrived) and queues the read-back of
        its loss; returns the step's index for ``result``."""
        i = self._run
        k = i % len(self.slots)
        g = self.slots[k]
        if self._ready[k] is not None:
            torch.cuda.current_stream().wait_event(self._ready[k])
            self._ready[k] = None
        loss = g.step()
        computed = torch.cuda.Event()
        computed.record()
        with torch.cuda.stream(self._out_stream):  # the read-back never sits between two replays on the compute stream
            self._out_stream.wait_event(computed)
            self._loss_host[k].copy_(loss.detach().reshape(-1), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._out_stream)
        self._done[k] = ev
        self._run += 1
        return i

    def result(self, i: int) -> float:
        """Host value of step i's loss; call it at most ``slots - 1`` launches late (the slot's word is reused)."""
        k = i % len(self.slots)
        self._done[k].synchronize()
        ws = self.slots[k]._ws
        if ws is not None and int(ws.status[1]) != 0:
            # the instance buffer was too small for this (or, with a shared workspace, the following) frame: its image
            # and gradients are incomplete.  The slots re-capture with a larger buffer at their next step.
            self.overflowed_steps.append(i)
            import warnings
            warnings.warn(f"riggs_b200: step {i} overflowed the instance buffer; its gradients are incomplete "
                          f"(PipelinedSteps.overflowed_steps) -- size the buffer with fused.measure_instances")
        v = self._loss_host[k]
        return float(v) if v.numel() == 1 else v.clone()

    @property
    def pose_grads(self) -> Dict[str, torch.Tensor]:
        return self.slots[(self._run - 1) % len(self.slots)].pose_grads


class RotatingSteps:
    """Training views resident in HBM, visited round-robin: one captured graph per group of V views (V = views this rank
    renders per optimiser step), all graphs working in the same workspace.  ``step()`` replays the next group's graph
    without copying any input -- camera, pose and target change every step, as they do in training, at no cost.  The
    instance buffer is sized once from the LARGEST instance count over all views (``fused.measure_instances``)."""

    def __init__(self, hp: HotPath, cams, poses, targets, views_per_step: int = 1, warmup: int = 2):
        from .fused import WorkspacePool
        self.hp = hp
        self.pool = WorkspacePool()
        V = int(views_per_step)
        dev = hp.device
        dposes = [{k: v.detach().to(dev) for k, v in p.items()} for p in poses]
        self.max_instances = hp.measure_instances([c.to(dev) for c in cams], dposes, self.pool)
        groups = [list(range(i, i + V)) for i in range(0, len(cams) - V + 1, V)]
        pick = lambda lst, g: [lst[i] for i in g] if V > 1 else lst[g[0]]
        self.slots = [GraphedStep(hp, pick(cams, g), pick(poses, g), pick(targets, g), warmup=warmup, pool=self.pool)
                      for g in groups]
        self.launches_per_replay = self.slots[0].launches_per_replay
        self._i = 0

    def step(self) -> torch.Tensor:
        g = self.slots[self._i % len(self.slots)]
        self._i += 1
        return g.step()

    @property
    def overflowed(self) -> bool:
        return self.slots[(self._i - 1) % len(self.slots)].overflowed


class PoseSequenceRenderer:
    """Forward-only rendering of a novel-pose frame sequence from one camera: the loop of the reference's
    ``render_rig.py:250-334`` (``generate_random_motion``: one ``skeleton.step`` + ``render`` per frame, poses from
    ``skeleton_utils/interpolation_utils.py:58-86``) with the per-frame work captured once as a CUDA graph.

    The whole pose sequence of this rank sits in device memory ((F,J,4) rotations, (F,3) translations); the graph
    itself picks frame ``idx`` out of it and advances ``idx``, so a frame costs one replay and no host -> device copy.
    ``slots`` graphs with their own output images alternate, so that frame f's image can travel to the host (or be
    consumed on another stream) while frame f+1 is rendered.  No gradient state is kept and no collective is involved:
    ranks render disjoint contiguous frame ranges (``dist.shard_frames``)."""

    def __init__(self, hp: HotPath, cam, n_joints: int, max_frames: int, slots: int = 2, warmup: int = 2):
        from .fused import WorkspacePool
        if not hp.fused:
            raise ValueError("PoseSequenceRenderer needs HotPath(fused=True)")
        self.hp = hp
        dev = hp.device
        self.cam = cam.to(dev) if hasattr(cam, "to") else cam
        self.rot = torch.zeros(max_frames, n_joints, 4, device=dev)
        self.rot[..., 0] = 1.0
        self.trans = torch.zeros(max_frames, 3, device=dev)
        self.idx = torch.zeros(1, dtype=torch.long, device=dev)
        self.n_frames = 0
        self.pool = WorkspacePool()
        self.images, self.graphs, self._free_ev = [], [], []
        self.launches_per_replay = 0
        self._warmup = max(1, warmup)
        self._slots = slots
        self._captured = False

    def _frame(self):
        hp = self.hp
        pose = {"local_rotation": self.rot.index_select(0, self.idx)[0], "global_trans": self.trans.index_select(0, self.idx)[0]}
        old, hp.pool = hp.pool, self.pool
        try:
            out = hp.forward(self.cam, pose)
        finally:
            hp.pool = old
        self.idx.add_(1)
        return out["render"]

    def load(self, rotations: torch.Tensor, translations: torch.Tensor) -> None:
        """Uploads a pose sequence ((F,J,4), (F,3); device or pinned host tensors) and rewinds to its first frame."""
        F = rotations.shape[0]
        if F > self.rot.shape[0]:
            raise ValueError(f"sequence of {F} frames exceeds the renderer's capacity of {self.rot.shape[0]}")
        self.rot[:F].copy_(rotations, non_blocking=True)
        self.trans[:F].copy_(translations, non_blocking=True)
        self.n_frames = F
        self.idx.zero_()
        if not self._captured:
            self._capture()

    def _capture(self):
        from . import _lib
        s = torch.cuda.Stream(device=self.hp.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s), torch.no_grad():
            # size the instance buffer over the sequence's poses (every 8th frame), not from its first frame alone
            fr = list(range(0, self.n_frames, 8))
            poses = [{"local_rotation": self.rot[f], "global_trans": self.trans[f]} for f in fr]
            self.max_instances = self.hp.measure_instances([self.cam] * len(fr), poses, self.pool)
            for _ in range(self._warmup):
                self._frame()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        for _ in range(self._slots):
            g = torch.cuda.CUDAGraph()
            n0 = _lib.lib().rg_launch_count()
            with torch.cuda.graph(g), torch.no_grad():
                img = self._frame()
            self.launches_per_replay = int(_lib.lib().rg_launch_count() - n0)
            self.graphs.append(g)
            self.images.append(img)
            self._free_ev.append(None)
        self.idx.zero_()
        self._captured = True
        self._n = 0
        ws = self.pool.free and next(iter(self.pool.free.values()))[0]
        self._ws = ws or None

    def rewind(self):
        self.idx.zero_()

    def render_next(self):
        """Renders the next frame of the loaded sequence; returns (slot, image) -- the image is the slot's static
        buffer, valid until the slot is replayed again (``release`` tells the renderer when its consumer is done)."""
        k = self._n % len(self.graphs)
        if self._free_ev[k] is not None:
            torch.cuda.current_stream().wait_event(self._free_ev[k])
            self._free_ev[k] = None
        self.graphs[k].replay()
        self._n += 1
        return k, self.images[k]

    def release(self, slot: int, event: "torch.cuda.Event") -> None:
        """The consumer of slot's image (e.g. a device -> host copy on another stream) finishes at `event`."""
        self._free_ev[slot] = event

    @property
    def overflowed(self) -> bool:
        torch.cuda.synchronize()
        return self._ws is not None and int(self._ws.status[1]) != 0
