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
                 fused: bool = True, lambda_dssim: float = 0.0):
        """fused=True: one C-ABI call per direction (rg_skinned_forward/backward, no host sync);
        fused=False: the reference's call sequence (deform_by_pose, then render()) on the same kernels."""
        self.fused = fused
        self.lambda_dssim = float(lambda_dssim)  # > 0: the trainer's L1 + SSIM image loss (train_rig.py:508-509)
        self.device = torch.device(device)
        self.pc = GaussianParams.from_scene(scene, device=device)
        self.warp = sw.SkeletonWarp(scene.joints, scene.parents, scene.node_radius_raw,
                                    use_skinning_weight_mlp=False, use_template_offsets=False, device=device)
        self.mlp_weight = scene.mlp_weight.to(device) if (use_mlp_weight and scene.mlp_weight is not None) else None
        self.offsets = scene.template_offsets.to(device) if (use_offsets and scene.template_offsets is not None) else None
        self.pipe = _Pipe()
        self.bg = torch.zeros(3, device=device)
        self.isotropic = self.pc.use_isotropic_gs
        self.flat = None
        self.pool = None  # fused.WorkspacePool override (GraphedStep keeps a private one)

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

    def forward(self, cam, pose: Dict[str, torch.Tensor]):
        """pose: {'local_rotation': (J,4), 'global_trans': (3,)} on the device (may require grad)."""
        if self.fused:
            from .fused import render_skinned
            return render_skinned(cam, self.pc, self.warp, pose, self.pipe, self.bg, mlp_weight=self.mlp_weight,
                                  offsets=self.offsets, grad_out=self.flat.views if self.flat is not None else None,
                                  pool=self.pool,
                                  reduce=self.flat if getattr(self.flat, "overlap", False) else None)
        joints = self.warp.nodes[:, :3]
        G, node_rot, d_nodes = sw.fk(pose["local_rotation"], pose["global_trans"], joints, self.warp.parents_i32)
        mask = self.pc._motion_mask.reshape(-1) if self.pc._motion_mask is not None else None
        d_xyz, d_rot, _ = sw.skin(self.pc.get_xyz, joints, self.warp.parents_i32, self.warp._node_radius, G, node_rot,
                                  pose["global_trans"], mask, self.mlp_weight, self.offsets, 0, False)
        if self.isotropic:  # the trainer zeroes d_rotation for isotropic Gaussians (train_rig.py:422-423)
            d_rot = 0.0
        return render(cam, self.pc, self.pipe, self.bg, d_xyz, d_rot, 0.0)

    def step(self, cam, pose, target: torch.Tensor) -> torch.Tensor:
        """Forward + backward of one view; returns the (device) loss.  Gradients accumulate in .grad."""
        out = self.forward(cam, pose)
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

    def __init__(self, hp: HotPath, cam, pose: Dict[str, torch.Tensor], target: torch.Tensor, warmup: int = 3,
                 pool=None):
        if not hp.fused:
            raise ValueError("GraphedStep needs HotPath(fused=True): the unfused path synchronises with the host")
        from .fused import WorkspacePool
        import copy
        self.hp = hp
        dev = hp.device
        self.pool = pool if pool is not None else WorkspacePool()  # a shared pool = one workspace for several graphs
        self.cam = copy.copy(cam)
        self.cam.world_view_transform = cam.world_view_transform.to(dev).clone()
        self.cam.full_proj_transform = cam.full_proj_transform.to(dev).clone()
        self.cam.camera_center = cam.camera_center.to(dev).clone()
        self.pose = {k: v.detach().to(dev).clone().requires_grad_(True) for k, v in pose.items()}
        self.target = target.to(dev).clone()
        self.loss: Optional[torch.Tensor] = None
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._event: Optional[torch.cuda.Event] = None
        self._warmup = max(int(warmup), 1)
        self._capture()

    def _one(self):
        hp = self.hp
        hp.zero_grad()
        for v in self.pose.values():
            v.grad = None
        old, hp.pool = getattr(hp, "pool", None), self.pool
        try:
            return hp.step(self.cam, self.pose, self.target)
        finally:
            hp.pool = old

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
        self._reduces_in_graph = bool(self.hp.flat is not None and self.hp.flat.reduced_in_backward)
        self._ws = self.pool.free and next(iter(self.pool.free.values()))[0]
        self._captured_binning = self._ws.binning.data_ptr() if self._ws else None
        self._event = None

    @property
    def pose_grads(self) -> Dict[str, torch.Tensor]:
        return {k: v.grad for k, v in self.pose.items()}

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
        if cam is not None:
            self.cam.world_view_transform.copy_(cam.world_view_transform, non_blocking=True)
            self.cam.full_proj_transform.copy_(cam.full_proj_transform, non_blocking=True)
            self.cam.camera_center.copy_(cam.camera_center, non_blocking=True)
        if pose is not None:
            with torch.no_grad():
                for k, v in pose.items():
                    self.pose[k].copy_(v, non_blocking=True)
        if target is not None:
            self.target.copy_(target, non_blocking=True)
        self.graph.replay()
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

    def __init__(self, hp: HotPath, cam, pose: Dict[str, torch.Tensor], target: torch.Tensor, warmup: int = 3,
                 slots: int = 2, share_workspace: bool = True):
        from .fused import WorkspacePool
        self.hp = hp
        # the slots run one after the other on one stream, so they can work in the same geometry / binning / image
        # buffers (and keep that state in L2 from step to step); only the inputs are per slot
        pool = WorkspacePool() if share_workspace else None
        self.slots = [GraphedStep(hp, cam, pose, target, warmup=warmup, pool=pool) for _ in range(slots)]
        self.launches_per_replay = self.slots[0].launches_per_replay
        self._copy_stream = torch.cuda.Stream(device=hp.device)
        self._out_stream = torch.cuda.Stream(device=hp.device)
        self._ready = [None] * slots   # inputs of the slot's next step have arrived
        self._done = [None] * slots    # the slot's last replay (and its loss copy) has finished
        self._loss_host = [torch.zeros(1).pin_memory() for _ in range(slots)]
        self._fed = 0
        self._run = 0

    def feed(self, cam=None, pose=None, target=None) -> None:
        """Queues the host -> device copies of the NEXT unfed step's inputs (pinned tensors make them asynchronous)."""
        k = self._fed % len(self.slots)
        g, cs = self.slots[k], self._copy_stream
        if self._done[k] is not None:
            cs.wait_event(self._done[k])  # the slot's previous step no longer reads these buffers
        with torch.cuda.stream(cs), torch.no_grad():
            if cam is not None:
                g.cam.world_view_transform.copy_(cam.world_view_transform, non_blocking=True)
                g.cam.full_proj_transform.copy_(cam.full_proj_transform, non_blocking=True)
                g.cam.camera_center.copy_(cam.camera_center, non_blocking=True)
            if pose is not None:
                for name, v in pose.items():
                    g.pose[name].copy_(v, non_blocking=True)
            if target is not None:
                g.target.copy_(target, non_blocking=True)
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
            self._loss_host[k].copy_(loss.detach().reshape(1), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._out_stream)
        self._done[k] = ev
        self._run += 1
        return i

    def result(self, i: int) -> float:
        """Host value of step i's loss; call it at most ``slots - 1`` launches late (the slot's word is reused)."""
        k = i % len(self.slots)
        self._done[k].synchronize()
        return float(self._loss_host[k])

    @property
    def pose_grads(self) -> Dict[str, torch.Tensor]:
        return self.slots[(self._run - 1) % len(self.slots)].pose_grads
