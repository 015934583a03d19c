"""Fused hot path: pose -> forward kinematics -> skinning -> activations -> rasterize, forward and backward,
as ONE C-ABI call per direction (rg_skinned_forward / rg_skinned_backward) with no host round trip.

``render_skinned`` is the opt-in companion of ``gaussian_renderer.render()`` (same output dict): instead of taking
``d_xyz / d_rotation / d_scaling`` that a separate ``SkeletonWarp.deform_by_pose`` call produced as N-sized
tensors, it takes the skeleton module and the pose and lets the kernels skin the Gaussians in registers.
The reference has no such function; what it replaces is the call sequence of train_rig.py:386-414 + :488
(``deform_gaussians`` then ``render``).

Workspaces (the reference's geomBuffer / binningBuffer / imgBuffer) are persistent and reused across frames.
The instance list is bounded by the binning buffer's capacity; ``num_rendered`` never travels to the host on the
critical path.  It is read back asynchronously and checked one call late; the first frame of a new
(P, width, height) is checked synchronously to size the buffer.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional

import torch

from . import _lib
from ._lib import check, f32c, ptr, stream_ptr


class _Workspace:
    def __init__(self, P, W, H, device, binning_bytes):
        L = _lib.lib()
        self.key = (P, W, H)
        self.geom = torch.empty(int(L.rg_geom_bytes(P, W, H)), dtype=torch.uint8, device=device)
        self.image = torch.empty(int(L.rg_image_bytes(W, H)), dtype=torch.uint8, device=device)
        self.binning = torch.empty(int(binning_bytes), dtype=torch.uint8, device=device)
        self.status = torch.zeros(2, dtype=torch.int32).pin_memory()
        self.status_event: Optional[torch.cuda.Event] = None
        self.c = _lib.RgWorkspace(self.geom.data_ptr(), self.geom.numel(), self.image.data_ptr(), self.image.numel(),
                                  self.binning.data_ptr(), self.binning.numel())

    def grow_binning(self, nbytes):
        self.binning = torch.empty(int(nbytes), dtype=torch.uint8, device=self.geom.device)
        self.c.binning = self.binning.data_ptr()
        self.c.binning_bytes = self.binning.numel()


class WorkspacePool:
    """Free list of workspaces per (P, W, H); one is held from a forward until its backward (or its release)."""

    def __init__(self, headroom: float = 1.5, initial_bytes: Optional[int] = None):
        self.free: Dict[tuple, list] = {}
        self.capacity_bytes: Dict[tuple, int] = {}
        self.headroom = headroom
        self.initial_bytes = initial_bytes  # binning bytes of a first-ever workspace (default: 8 tiles per Gaussian)

    def acquire(self, P, W, H, device):
        key = (P, W, H, str(device))
        lst = self.free.setdefault(key, [])
        if lst:
            return lst.pop(), False
        nbytes = self.capacity_bytes.get(key, 0)
        fresh = nbytes == 0
        if fresh:
            nbytes = self.initial_bytes or int(_lib.lib().rg_binning_bytes(max(8 * P, 1 << 18)))
        return _Workspace(P, W, H, device, nbytes), fresh

    def release(self, ws, device):
        P, W, H = ws.key
        self.free.setdefault((P, W, H, str(device)), []).append(ws)

    def reserve(self, P, W, H, device, num_rendered: int):
        """Sizes the binning buffers of this shape for `num_rendered` instances (times the headroom): call it with
        the largest count seen over a pass of the training views (``measure_instances``) so that no later view
        overflows a buffer that was sized from the first frame alone."""
        key = (P, W, H, str(device))
        want = int(_lib.lib().rg_binning_bytes(int(num_rendered * self.headroom) + 256))
        if want > self.capacity_bytes.get(key, 0):
            self.capacity_bytes[key] = want
            for ws in self.free.get(key, []):
                if ws.binning.numel() < want:
                    ws.grow_binning(want)

    def note_capacity(self, ws, device):
        P, W, H = ws.key
        key = (P, W, H, str(device))
        self.capacity_bytes[key] = max(self.capacity_bytes.get(key, 0), ws.binning.numel())


_POOL = WorkspacePool()
_SCREENSPACE_ZEROS: Dict[tuple, torch.Tensor] = {}
_SIDE: Dict[str, torch.cuda.Stream] = {}


def _side_stream(dev) -> torch.cuda.Stream:
    s = _SIDE.get(str(dev))
    if s is None:
        s = _SIDE[str(dev)] = torch.cuda.Stream(device=dev)
    return s


class _Lease:
    """Returns the workspace to its pool after the backward -- or when autograd drops the graph without one."""

    def __init__(self, pool, ws, dev):
        self.pool, self.ws, self.dev = pool, ws, dev

    def release(self):
        if self.ws is not None:
            self.pool.release(self.ws, self.dev)
            self.ws = None

    def __del__(self):
        self.release()


def _scene_struct(P, J, sh_degree, M_rest, isotropic, zero_d_rot, W, H, tfx, tfy, scale_mod, t, blend_mode=0, split=None):
    s = _lib.RgSkinnedScene()
    s.blend_mode = int(blend_mode)
    if split is not None and split.count > 1:
        s.split_index, s.split_count = split.index, split.count
    s.P, s.J, s.sh_degree, s.M_rest = P, J, sh_degree, M_rest
    s.isotropic, s.zero_d_rotation, s.width, s.height = int(isotropic), int(zero_d_rot), W, H
    s.tan_fovx, s.tan_fovy, s.scale_modifier = tfx, tfy, scale_mod
    for name, ten in t.items():
        setattr(s, name, ptr(ten))
    return s


def _grad_acc_view(ws, P, W, H) -> torch.Tensor:
    """The 10 P floats of per-Gaussian screen-space accumulators inside a workspace's geometry buffer (rg_state_ptrs)."""
    out = (C.c_void_p * 13)()
    _lib.lib().rg_state_ptrs(P, 0, W, H, ws.geom.data_ptr(), None, ws.image.data_ptr(), out)
    off = int(out[12]) - ws.geom.data_ptr()
    return ws.geom[off:off + 40 * P].view(torch.float32)


class _SkinnedRender(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, scaling_raw, rotation_raw, opacity_raw, features_dc, features_rest, node_radius_raw,
                local_rot, global_trans, motion_mask, mlp_weight, offsets, screenspace_points, cfg):
        if not xyz.is_cuda:
            raise _lib.RiggsError("render_skinned needs CUDA tensors; riggs_b200 has no CPU path")
        dev = xyz.device
        P, J = xyz.shape[0], cfg["joints"].shape[0]
        W, H = cfg["width"], cfg["height"]
        t = dict(xyz=f32c(xyz), scaling_raw=f32c(scaling_raw), rotation_raw=f32c(rotation_raw),
                 opacity_raw=f32c(opacity_raw), features_dc=f32c(features_dc), features_rest=f32c(features_rest),
                 motion_mask=f32c(motion_mask), mlp_weight=f32c(mlp_weight), offsets=f32c(offsets),
                 joints=cfg["joints"], parents=cfg["parents_i32"], node_radius_raw=f32c(node_radius_raw),
                 local_rot=f32c(local_rot), global_trans=f32c(global_trans), viewmatrix=cfg["viewmatrix"],
                 projmatrix=cfg["projmatrix"], campos=cfg["campos"], background=cfg["bg"])
        isotropic = scaling_raw.shape[-1] == 1
        M_rest = features_rest.shape[1]
        scene = _scene_struct(P, J, cfg["sh_degree"], M_rest, isotropic, cfg["zero_d_rotation"], W, H, cfg["tanfovx"],
                              cfg["tanfovy"], cfg["scale_modifier"], t, cfg.get("blend_mode", 0), cfg.get("split"))
        split = cfg.get("split")
        if split is not None and split.count > 1:
            # one view over several GPUs (dist.ViewSplit): this rank fills its tile rows of a zeroed image ...
            if (split.P, split.width, split.height) != (P, W, H):
                raise _lib.RiggsError("render_skinned: the ViewSplit was built for another (P, width, height)")
            split.images.buf.zero_()
            color, depth, alpha = split.image_views()
        else:
            split = None
            color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
            depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
            alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        d_nodes = torch.empty((J, 3), dtype=torch.float32, device=dev)
        dbg = cfg.get("debug_outputs")
        L = _lib.lib()
        pool = cfg.get("pool") or _POOL
        ws, fresh = pool.acquire(P, W, H, dev)
        capturing = torch.cuda.is_current_stream_capturing()
        if capturing and fresh:
            raise _lib.RiggsError("render_skinned under CUDA-graph capture needs a workspace sized by an eager call "
                                  "of the same shape first")
        with torch.cuda.device(dev):
            # late check of the previous frame that used this workspace
            if not capturing and ws.status_event is not None and ws.status_event.query() and int(ws.status[1]) != 0:
                import warnings
                warnings.warn("riggs_b200: the previous frame overflowed the instance buffer and was rendered "
                              "incompletely; growing the buffer")
                ws.status_event = None
                fresh = True  # fall through to the synchronous sizing path
            while True:
                check(L.rg_skinned_forward(C.byref(scene), C.byref(ws.c), ptr(color), ptr(depth), ptr(alpha), ptr(radii),
                                           ptr(d_nodes), ptr(dbg["scales"]) if dbg else None,
                                           ptr(dbg["rotations"]) if dbg else None,
                                           ptr(dbg["opacities"]) if dbg else None, stream_ptr()), "rg_skinned_forward")
                check(L.rg_skinned_status(C.byref(scene), C.byref(ws.c), ws.status.data_ptr(), stream_ptr()),
                      "rg_skinned_status")
                if capturing:  # the owner of the graph checks ws.status after its replays
                    ws.status_event = None
                    break
                if not fresh:
                    ws.status_event = torch.cuda.Event()
                    ws.status_event.record()
                    break
                torch.cuda.current_stream().synchronize()  # first frame of this shape only
                R, overflow = int(ws.status[0]), int(ws.status[1])
                want = int(_lib.lib().rg_binning_bytes(int(R * pool.headroom) + 256))
                if overflow or ws.binning.numel() < want:
                    # too small (or no headroom for the coming frames): grow and render this first frame again,
                    # so the instance list the backward will read lives in the buffer the workspace keeps
                    ws.grow_binning(want)
                    pool.note_capacity(ws, dev)
                    continue
                pool.note_capacity(ws, dev)
                break
        if split is not None:
            # ... and the sum over the ranks' disjoint rows is the whole image, on every rank
            with torch.cuda.device(dev):
                split.sum_images()
            color, depth, alpha = color.clone(), depth.clone(), alpha.clone()
        ctx.split = split
        ctx.set_materialize_grads(False)  # outputs the loss does not use arrive as None, not as zero-filled images
        ctx.scene_tensors = t  # keeps the contiguous copies alive
        ctx.scene = scene
        ctx.ws = ws
        ctx.lease = _Lease(pool, ws, dev)
        ctx.dev = dev
        ctx.grad_out = cfg.get("grad_out")
        ctx.reduce = cfg.get("reduce")
        ctx.accumulate = bool(cfg.get("accumulate"))
        if ctx.accumulate and not ctx.grad_out:
            raise _lib.RiggsError("render_skinned(accumulate=True) adds into caller-owned buffers: pass grad_out")
        if not any(ctx.needs_input_grad):
            ctx.lease.release()  # inference: the state is not needed any more
        ctx.shapes = (P, J, W, H, M_rest, isotropic)
        ctx.save_for_backward(radii, alpha)
        ctx.mark_non_differentiable(radii)
        ctx.num_rendered_host = ws.status  # valid after the next synchronisation
        cfg["_status"] = ws.status
        return color, depth, alpha, radii, d_nodes

    @staticmethod
    def backward(ctx, g_color, g_depth, g_alpha, _g_radii, g_nodes):
        if ctx.ws is None:
            # the forward state lives in a pooled workspace that went back to the pool after the first backward
            raise _lib.RiggsError("render_skinned: backward was already run for this frame (the workspace holding its "
                                  "forward state has been released); render the frame again for a second backward")
        radii, alpha = ctx.saved_tensors
        P, J, W, H, M_rest, isotropic = ctx.shapes
        dev = ctx.dev
        t = ctx.scene_tensors
        opts = dict(dtype=torch.float32, device=dev)
        z = lambda *s: torch.zeros(s, **opts)
        g_color = f32c(g_color) if g_color is not None else z(3, H, W)
        g_depth = f32c(g_depth) if g_depth is not None else None  # NULL = no cotangent (rg_skinned_backward)
        g_alpha = f32c(g_alpha) if g_alpha is not None else None
        g_nodes = f32c(g_nodes) if g_nodes is not None else None
        go = ctx.grad_out or {}

        def out(name, shape):
            """Gradient buffer: the caller's (e.g. a view of the flat all-reduce buffer) or a fresh one."""
            b = go.get(name)
            if b is not None:
                if tuple(b.shape) != tuple(shape) or b.dtype != torch.float32 or not b.is_contiguous() or b.device != dev:
                    raise _lib.RiggsError(f"grad_out[{name!r}] must be a contiguous fp32 tensor of shape {tuple(shape)}")
                return b
            return torch.empty(shape, **opts)

        d_xyz = out("xyz", (P, 3))
        d_scaling = out("scaling", (P, 1 if isotropic else 3))
        d_rotation = out("rotation", (P, 4))
        d_opacity = out("opacity", (P, 1))
        d_fdc = out("features_dc", (P, 1, 3))
        d_frest = out("features_rest", (P, M_rest, 3))
        d_mean2D = torch.empty((P, 3), **opts)
        d_local_rot = torch.empty((J, 4), **opts)
        d_trans = torch.empty((3,), **opts)
        d_radius = out("node_radius", (J,))
        need = ctx.needs_input_grad
        d_mask = torch.empty((P,), **opts) if (t["motion_mask"] is not None and need[9]) else None
        d_mlp = torch.empty((P, J - 1), **opts) if (t["mlp_weight"] is not None and need[10]) else None
        d_off = torch.empty((P, 3), **opts) if (t["offsets"] is not None and need[11]) else None
        red = ctx.reduce
        args = (ptr(radii), ptr(alpha), ptr(g_color), ptr(g_depth), ptr(g_alpha),
                ptr(g_nodes), ptr(d_xyz), ptr(d_scaling), ptr(d_rotation), ptr(d_opacity), ptr(d_fdc), ptr(d_frest),
                ptr(d_mean2D), ptr(d_local_rot), ptr(d_trans), ptr(d_radius), ptr(d_mask), ptr(d_mlp), ptr(d_off))
        L = _lib.lib()
        with torch.cuda.device(dev):
            if ctx.split is not None:
                # blend backward over this rank's tile rows, sum of the per-Gaussian accumulators over the ranks, then the
                # per-Gaussian and skeleton backward replicated: every rank ends with the full gradients
                acc = 8 if ctx.accumulate else 0
                check(L.rg_skinned_backward_part(C.byref(ctx.scene), C.byref(ctx.ws.c), 1, 0, 0, *args, stream_ptr()),
                      "rg_skinned_backward_part")
                ctx.split.sum_grad2d(_grad_acc_view(ctx.ws, P, W, H))
                check(L.rg_skinned_backward_part(C.byref(ctx.scene), C.byref(ctx.ws.c), 6 | acc, 0, P, *args, stream_ptr()),
                      "rg_skinned_backward_part")
            elif red is not None and red.can_overlap(P) and all(n in go for n in red.names):
                # Data parallel: the parameter gradients go straight into the ranks' peer-mapped flat buffer, and the
                # rows of chunk c are summed over the ranks on a second stream while chunk c+1 is computed here.
                acc = 8 if ctx.accumulate else 0
                part = lambda phases, r0, r1: check(L.rg_skinned_backward_part(
                    C.byref(ctx.scene), C.byref(ctx.ws.c), phases | (acc if phases & 6 else 0), r0, r1, *args, stream_ptr()),
                    "rg_skinned_backward_part")
                main = torch.cuda.current_stream()
                side = _side_stream(dev)
                chunks = red.row_chunks(P)
                part(1, 0, 0)
                for c, (r0, r1) in enumerate(chunks):
                    last = c == len(chunks) - 1
                    part(2, r0, r1)
                    if last:
                        part(4, 0, 0)  # the skeleton gradients travel with the last chunk
                    ev = torch.cuda.Event()
                    ev.record(main)
                    with torch.cuda.stream(side):
                        side.wait_event(ev)
                        red.all_reduce_rows(P, r0, r1, with_rest=last, barrier_after=last)
                main.wait_stream(side)
            elif ctx.accumulate:  # a later view of a multi-view batch: add into the parameter-gradient buffers
                check(L.rg_skinned_backward_part(C.byref(ctx.scene), C.byref(ctx.ws.c), 7 | 8, 0, P, *args, stream_ptr()),
                      "rg_skinned_backward_part")
            else:
                check(L.rg_skinned_backward(C.byref(ctx.scene), C.byref(ctx.ws.c), *args, stream_ptr()), "rg_skinned_backward")
        ctx.lease.release()
        ctx.ws = None
        if d_mask is not None:
            d_mask = d_mask.view_as(t["motion_mask"])
        # A gradient that was written into a caller-provided buffer is delivered through that buffer only (autograd
        # would otherwise clone it into param.grad: one more pass over ~240 B per Gaussian).
        g = lambda name, ten: None if name in go else ten
        return (g("xyz", d_xyz), g("scaling", d_scaling), g("rotation", d_rotation), g("opacity", d_opacity),
                g("features_dc", d_fdc), g("features_rest", d_frest), g("node_radius", d_radius), d_local_rot, d_trans,
                d_mask, d_mlp, d_off, d_mean2D, None)


class _RenderResult(dict):
    """The result dict of ``render()``; ``visibility_filter`` (= ``radii > 0``, one more pass over the P radii) is computed
    when it is first asked for -- a training step that does not densify never asks."""

    def __missing__(self, key):
        if key == "visibility_filter":
            v = self["radii"] > 0
            self[key] = v
            return v
        raise KeyError(key)

    def get(self, key, default=None):
        try:
            return self[key]
        except KeyError:
            return default

    def __contains__(self, key):
        return key == "visibility_filter" or dict.__contains__(self, key)


def render_skinned(viewpoint_camera, pc, skeleton, pose: Dict[str, torch.Tensor], pipe, bg_color: torch.Tensor,
                   scaling_modifier=1.0, mlp_weight=None, offsets=None, zero_d_rotation: Optional[bool] = None,
                   debug_outputs=None, pool: Optional[WorkspacePool] = None, grad_out=None, reduce=None,
                   accumulate: bool = False, split=None):
    """Fused deform + render.  ``pc``: object with the GaussianModel raw tensors (_xyz, _scaling, _rotation, _opacity,
    _features_dc, _features_rest, motion_mask, active_sh_degree); ``skeleton``: riggs_b200 SkeletonWarp (joints,
    parents, _node_radius); ``pose``: {'local_rotation': (J,4), 'global_trans': (3,)}.
    ``grad_out``: optional dict of preallocated gradient buffers (keys xyz, scaling, rotation, opacity, features_dc,
    features_rest, node_radius) the backward kernels write into directly -- e.g. the views of ``dist.FlatGrads``.
    ``reduce``: that ``dist.FlatGrads`` itself; when its buffer is peer-mapped the backward sums the gradient over the
    ranks chunk by chunk while it is still computing later chunks (``FlatGrads.all_reduce`` then has nothing left to do).
    ``accumulate``: the backward ADDS the parameter gradients into ``grad_out`` instead of overwriting them -- the second
    and later views of a multi-view batch (rg_skinned_backward_part, phase bit 8).
    ``split``: a ``dist.ViewSplit``: this ONE view is rendered by all the ranks together (each bins and blends every
    count-th tile row; the image and the gradients come out complete and identical on every rank).
    Returns the dict of ``render()`` plus ``d_nodes`` and ``status``: a pinned host tensor {num_rendered, overflow}
    that is valid once the stream has been synchronised (``frame_overflowed``)."""
    xyz = pc._xyz
    # the reference's dummy leaf whose .grad receives dL/dmean2D (gaussian_renderer/__init__.py:45-49); its values are
    # never read, so every call wraps the same zero buffer in a new leaf instead of filling a fresh one
    key = (xyz.shape[0], str(xyz.device))
    zeros = _SCREENSPACE_ZEROS.get(key)
    if zeros is None:
        _SCREENSPACE_ZEROS.clear()
        zeros = _SCREENSPACE_ZEROS[key] = torch.zeros_like(xyz, requires_grad=False)
    screenspace_points = zeros.detach().requires_grad_(True)
    isotropic = pc._scaling.shape[-1] == 1
    cfg = dict(joints=f32c(skeleton.nodes[:, :3].detach()), parents_i32=skeleton.parents_i32,
               width=int(viewpoint_camera.image_width), height=int(viewpoint_camera.image_height),
               tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5),
               scale_modifier=float(scaling_modifier), sh_degree=int(pc.active_sh_degree),
               viewmatrix=f32c(viewpoint_camera.world_view_transform), projmatrix=f32c(viewpoint_camera.full_proj_transform),
               campos=f32c(viewpoint_camera.camera_center), bg=f32c(bg_color),
               zero_d_rotation=isotropic if zero_d_rotation is None else zero_d_rotation, debug_outputs=debug_outputs,
               pool=pool, grad_out=grad_out, reduce=reduce, accumulate=accumulate, split=split,
               blend_mode=getattr(skeleton, "blend_mode", 0))
    mm = getattr(pc, "_motion_mask", None)
    mask = mm.reshape(-1) if mm is not None else None
    color, depth, alpha, radii, d_nodes = _SkinnedRender.apply(
        xyz, pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest, skeleton._node_radius,
        pose["local_rotation"], pose["global_trans"], mask, mlp_weight, offsets, screenspace_points, cfg)
    return _RenderResult({"render": color, "viewspace_points": screenspace_points, "radii": radii, "depth": depth,
                          "alpha": alpha, "bg_color": bg_color, "d_nodes": d_nodes, "status": cfg.get("_status")})


def frame_overflowed(out) -> bool:
    """True when the frame `out` (a ``render_skinned`` result) produced more (Gaussian, tile) instances than the
    workspace's instance buffer holds: the image and the gradients of that frame are incomplete and must not reach the
    optimiser.  Synchronises the current stream (call it where the loss is read on the host anyway).  The next
    ``render_skinned`` of the same shape grows the buffer; render the frame again."""
    torch.cuda.current_stream().synchronize()
    st = out.get("status")
    return st is not None and int(st[1]) != 0


def measure_instances(viewpoint_cameras, pc, skeleton, poses, pipe, bg_color, pool: Optional[WorkspacePool] = None,
                      **kw) -> int:
    """Renders every (camera, pose) pair once without gradients, reads each frame's instance count on the host and sizes
    the pool's instance buffers for the LARGEST one (``WorkspacePool.reserve``).  Returns that count.  Run it once over
    the training views before the first step (and after densification): later frames then cannot overflow a buffer that
    the first frame alone would have sized."""
    pool = pool or _POOL
    r_max = 0
    xyz = pc._xyz
    with torch.no_grad():
        for cam, pose in zip(viewpoint_cameras, poses):
            out = render_skinned(cam, pc, skeleton, pose, pipe, bg_color, pool=pool, **kw)
            torch.cuda.current_stream().synchronize()
            if int(out["status"][1]) != 0:  # grown by the next call; the count is valid even for an overflowed frame
                pool.reserve(xyz.shape[0], int(cam.image_width), int(cam.image_height), xyz.device, int(out["status"][0]))
            r_max = max(r_max, int(out["status"][0]))
    if viewpoint_cameras:
        cam = viewpoint_cameras[0]
        pool.reserve(xyz.shape[0], int(cam.image_width), int(cam.image_height), xyz.device, r_max)
    return r_max
