"""Optimiser step and densification statistics of the reference's training iteration on the B200 kernels
(csrc/optim.cu).

``FusedAdam`` takes the reference's own parameter-group list (scene/gaussian_model.py:203-217: dicts with 'params', 'lr',
'name') and does ``torch.optim.Adam(l, lr=0.0, eps=1e-15).step()`` for all groups in one kernel launch.  ``param_groups``
stays a list of dicts, so ``GaussianModel.update_learning_rate`` (:220-226) keeps working unchanged.  No CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from . import _lib
from ._lib import check, ptr, stream_ptr


class FusedAdam:
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8):
        groups = list(params)
        if groups and not isinstance(groups[0], dict):
            groups = [{"params": groups}]
        self.param_groups: List[dict] = []
        for g in groups:
            g = dict(g)
            g["params"] = list(g["params"])
            g.setdefault("lr", lr)
            self.param_groups.append(g)
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.state: Dict[torch.Tensor, dict] = {}
        self._step = 0
        n = sum(len(g["params"]) for g in self.param_groups)
        if n > 16:
            raise ValueError("FusedAdam: at most 16 parameter tensors")

    def zero_grad(self, set_to_none: bool = True):
        for g in self.param_groups:
            for p in g["params"]:
                if set_to_none:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.zero_()

    @torch.no_grad()
    def step(self):
        # torch.optim.Adam keeps state['step'] per parameter and advances it only for parameters that have a gradient:
        # entries are grouped by their own step count (one launch per distinct count -- one, unless some parameter
        # missed gradients on earlier iterations)
        buckets: Dict[int, list] = {}
        keep = []  # contiguous gradient copies must outlive the launch
        dev = None
        for g in self.param_groups:
            for p in g["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise _lib.RiggsError("FusedAdam needs contiguous fp32 CUDA parameters; riggs_b200 has no CPU path")
                st = self.state.get(p)
                if st is None:
                    st = self.state[p] = {"step": 0, "exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p)}
                st["step"] = int(st.get("step", 0)) + 1
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                keep.append(grad)
                dev = p.device
                buckets.setdefault(st["step"], []).append(
                    _lib.RgAdamGroup(ptr(p), ptr(grad), ptr(st["exp_avg"]), ptr(st["exp_avg_sq"]), p.numel(), float(g["lr"])))
        if not buckets:
            return
        self._step += 1
        with torch.cuda.device(dev):
            for step, entries in buckets.items():
                for i in range(0, len(entries), 16):  # rg_adam_step takes up to 16 groups per launch
                    part = entries[i:i + 16]
                    arr = (_lib.RgAdamGroup * len(part))(*part)
                    check(_lib.lib().rg_adam_step(len(part), C.cast(arr, C.c_void_p), step, self.betas[0], self.betas[1],
                                                  self.eps, stream_ptr()), "rg_adam_step")


@torch.no_grad()
def add_densification_stats(viewspace_point_tensor: torch.Tensor, radii: torch.Tensor, xyz_gradient_accum: torch.Tensor,
                            denom: torch.Tensor, max_radii2D: Optional[torch.Tensor] = None):
    """GaussianModel.add_densification_stats(viewspace_point_tensor, radii > 0) and, when max_radii2D is given, the
    max_radii2D[visibility_filter] = max(...) update of train_rig.py:335, in one launch.  ``radii`` int32 (P)."""
    g = viewspace_point_tensor.grad if viewspace_point_tensor.grad is not None else viewspace_point_tensor
    if not g.is_cuda:
        raise _lib.RiggsError("add_densification_stats needs CUDA tensors; riggs_b200 has no CPU path")
    g = g if g.is_contiguous() else g.contiguous()
    P = g.shape[0]
    if radii.dtype != torch.int32:
        radii = radii.to(torch.int32)
    for t in (xyz_gradient_accum, denom) + ((max_radii2D,) if max_radii2D is not None else ()):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P:
            raise ValueError("densification statistics must be contiguous fp32 tensors with one entry per Gaussian")
    with torch.cuda.device(g.device):
        check(_lib.lib().rg_densification_stats(P, ptr(g), ptr(radii), ptr(xyz_gradient_accum), ptr(denom),
                                                ptr(max_radii2D), stream_ptr()), "rg_densification_stats")


_COMPACT_SCRATCH: Dict[tuple, torch.Tensor] = {}


@torch.no_grad()
def compact_rows(keep: torch.Tensor, tensors: List[torch.Tensor]) -> List[torch.Tensor]:
    """``[t[keep] for t in tensors]`` for tensors that share their first dimension, with one scan of the mask and one
    copy kernel for all of them (csrc/compact.cu).  keep: bool (P,); tensors: contiguous, 4-byte element types."""
    if not keep.is_cuda:
        raise _lib.RiggsError("compact_rows needs CUDA tensors; riggs_b200 has no CPU path")
    P = keep.numel()
    keep_u8 = keep.reshape(-1).to(torch.uint8).contiguous()
    srcs = []
    for t in tensors:
        if t.shape[0] != P or t.element_size() != 4:
            raise ValueError("compact_rows: every tensor needs P rows of 4-byte elements")
        srcs.append(t if t.is_contiguous() else t.contiguous())
    L = _lib.lib()
    dev = keep.device
    nbytes = int(L.rg_compact_scratch_bytes(P))
    key = (str(dev), torch.cuda.current_stream(dev).cuda_stream, nbytes)
    sc = _COMPACT_SCRATCH.get(key)
    if sc is None:
        sc = _COMPACT_SCRATCH[key] = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    n = int(keep_u8.sum()) if P else 0  # the one host round trip, as with boolean indexing
    count = torch.empty((), dtype=torch.int32, device=dev)
    dsts = [torch.empty((n,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev) for t in srcs]
    if n:
        arr = (_lib.RgRowTensor * len(srcs))(*[_lib.RgRowTensor(ptr(s), ptr(d), (s.numel() // P) * 4)
                                              for s, d in zip(srcs, dsts)])
        with torch.cuda.device(dev):
            check(L.rg_compact_rows(P, ptr(keep_u8), len(srcs), C.cast(arr, C.c_void_p), ptr(sc), sc.numel(), ptr(count),
                                    stream_ptr()), "rg_compact_rows")
    return dsts


def prune_points(optimizer: FusedAdam, mask: torch.Tensor, stats: Optional[Dict[str, torch.Tensor]] = None):
    """GaussianModel.prune_points(mask) + _prune_optimizer (scene/gaussian_model.py:356-391) for a ``FusedAdam`` with
    one parameter per group: drops the rows where ``mask`` is True from every parameter, its Adam moments and the
    statistics in ``stats`` (xyz_gradient_accum, denom, max_radii2D ...).  Returns ({group name: new nn.Parameter},
    {stat name: new tensor}); the optimizer's groups and state are updated in place like the reference does."""
    keep = ~mask.reshape(-1).bool()
    params = [g["params"][0] for g in optimizer.param_groups]
    tensors: List[torch.Tensor] = [p.detach() for p in params]
    has_state = [optimizer.state.get(p) for p in params]
    for st in has_state:
        if st is not None:
            tensors += [st["exp_avg"], st["exp_avg_sq"]]
    stat_names = list(stats or {})
    tensors += [stats[k] for k in stat_names]
    out = compact_rows(keep, tensors)
    new_params: Dict[str, torch.nn.Parameter] = {}
    k = len(params)
    for i, (g, p, st) in enumerate(zip(optimizer.param_groups, params, has_state)):
        newp = torch.nn.Parameter(out[i].requires_grad_(True))
        if st is not None:
            del optimizer.state[p]
            optimizer.state[newp] = {"step": int(st.get("step", 0)), "exp_avg": out[k], "exp_avg_sq": out[k + 1]}
            k += 2
        g["params"][0] = newp
        new_params[g.get("name", str(i))] = newp
    return new_params, {n: out[k + j] for j, n in enumerate(stat_names)}


# ---- densification: clone / split / prune (scene/gaussian_model.py:393-505) -------------------------------------
# The reference appends the new points with a torch.cat per parameter and per Adam moment (cat_tensors_to_optimizer),
# then -- for a split -- removes the parents with a boolean index per tensor again (prune_points).  Here every step is
# ONE index list (kept rows, then the parents of the new rows) and two gather launches (csrc/compact.cu: the
# parameters, and the Adam moments with zeros for the new rows); a split never materialises the parents + children set.

def _groups(optimizer: FusedAdam):
    return {g.get("name", str(i)): g for i, g in enumerate(optimizer.param_groups)}


@torch.no_grad()
def regroup_rows(optimizer: FusedAdam, src_row: torch.Tensor, n_old_rows: int) -> Dict[str, torch.nn.Parameter]:
    """New parameters with ``len(src_row)`` rows: row r is a copy of old row ``src_row[r]``.  The Adam moments follow
    for the first ``n_old_rows`` entries of ``src_row`` (rows that survive) and start at zero for the rest (new
    points), as cat_tensors_to_optimizer does (:393-415).  Groups and optimizer state are replaced in place."""
    dev = src_row.device
    if not src_row.is_cuda:
        raise _lib.RiggsError("regroup_rows needs CUDA tensors; riggs_b200 has no CPU path")
    src_row = src_row.to(torch.int32).contiguous()
    n_out = src_row.numel()
    src_state = src_row.clone()
    src_state[n_old_rows:] = -1
    params = [g["params"][0] for g in optimizer.param_groups]
    P = params[0].shape[0]
    new = [torch.empty((n_out,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev) for p in params]
    states = [optimizer.state.get(p) for p in params]
    m_src, m_dst = [], []
    for p, st in zip(params, states):
        if st is not None:
            for k in ("exp_avg", "exp_avg_sq"):
                m_src.append(st[k])
                m_dst.append(torch.empty((n_out,) + tuple(p.shape[1:]), dtype=p.dtype, device=dev))
    L = _lib.lib()

    def gather(srcs, dsts, rows):
        if not srcs or n_out == 0:
            return
        arr = (_lib.RgRowTensor * len(srcs))(*[_lib.RgRowTensor(ptr(s.detach().contiguous()), ptr(d), (s.numel() // P) * 4)
                                              for s, d in zip(srcs, dsts)])
        check(L.rg_gather_rows(n_out, ptr(rows), len(srcs), C.cast(arr, C.c_void_p), stream_ptr()), "rg_gather_rows")

    with torch.cuda.device(dev):
        gather(params, new, src_row)
        gather(m_src, m_dst, src_state)
    out: Dict[str, torch.nn.Parameter] = {}
    k = 0
    for i, (g, p, st) in enumerate(zip(optimizer.param_groups, params, states)):
        newp = torch.nn.Parameter(new[i].requires_grad_(True))
        if st is not None:
            del optimizer.state[p]
            optimizer.state[newp] = {"step": int(st.get("step", 0)), "exp_avg": m_dst[k], "exp_avg_sq": m_dst[k + 1]}
            k += 2
        g["params"][0] = newp
        out[g.get("name", str(i))] = newp
    return out


def _reset_stats(stats: Optional[Dict[str, torch.Tensor]], n: int, device) -> Dict[str, torch.Tensor]:
    """densification_postfix (:437-439): the statistics start over at the new size."""
    return {k: torch.zeros((n,) + tuple(v.shape[1:]), dtype=v.dtype, device=device) for k, v in (stats or {}).items()}


def _scaling_activated(optimizer: FusedAdam) -> torch.Tensor:
    s = _groups(optimizer)["scaling"]["params"][0].detach()
    if s.shape[1] == 1:  # isotropic Gaussians keep one column (get_scaling, :105-110)
        s = s.repeat(1, 3)
    return torch.exp(s)


def build_rotation(r: torch.Tensor) -> torch.Tensor:
    """utils/general_utils.py:137-158."""
    q = r / torch.sqrt((r * r).sum(1))[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)


@torch.no_grad()
def densify_and_clone(optimizer: FusedAdam, stats, grads=None, grad_threshold=None, scene_extent=None,
                      percent_dense: float = 0.01, selected_pts_mask=None):
    """GaussianModel.densify_and_clone (:479-499): small Gaussians with a large view-space gradient are duplicated.
    Returns ({group name: new parameter}, fresh statistics)."""
    params = _groups(optimizer)
    P = params["xyz"]["params"][0].shape[0]
    if selected_pts_mask is None:
        selected_pts_mask = torch.norm(grads, dim=-1) >= grad_threshold
        selected_pts_mask = torch.logical_and(selected_pts_mask,
                                              _scaling_activated(optimizer).max(dim=1).values <= percent_dense * scene_extent)
    dev = selected_pts_mask.device
    sel = selected_pts_mask.reshape(-1).nonzero().reshape(-1).to(torch.int32)
    src = torch.cat([torch.arange(P, dtype=torch.int32, device=dev), sel])
    new = regroup_rows(optimizer, src, P)
    return new, _reset_stats(stats, src.numel(), dev)


@torch.no_grad()
def densify_and_split(optimizer: FusedAdam, stats, grads=None, grad_threshold=None, scene_extent=None,
                      percent_dense: float = 0.01, N: int = 2, selected_pts_mask=None, without_prune: bool = False,
                      samples: Optional[torch.Tensor] = None, generator: Optional[torch.Generator] = None):
    """GaussianModel.densify_and_split (:441-477): large Gaussians with a large gradient are replaced by N smaller
    ones drawn from them.  ``samples``: optional (N * n_selected, 3) standard-normal draws (the reference calls
    torch.normal(0, stds); here ``samples * stds``), so that a caller or a test controls the random numbers.
    Returns ({group name: new parameter}, fresh statistics)."""
    params = _groups(optimizer)
    P = params["xyz"]["params"][0].shape[0]
    dev = params["xyz"]["params"][0].device
    if selected_pts_mask is None:
        padded_grad = torch.zeros(P, device=dev)
        padded_grad[:grads.shape[0]] = grads.squeeze()
        selected_pts_mask = padded_grad >= grad_threshold
        selected_pts_mask = torch.logical_and(selected_pts_mask,
                                              _scaling_activated(optimizer).max(dim=1).values > percent_dense * scene_extent)
    selected_pts_mask = selected_pts_mask.reshape(-1)
    sel = selected_pts_mask.nonzero().reshape(-1).to(torch.int32)
    kept = (torch.arange(P, dtype=torch.int32, device=dev) if without_prune
            else (~selected_pts_mask).nonzero().reshape(-1).to(torch.int32))
    n_keep, n_new = kept.numel(), sel.numel() * N
    src = torch.cat([kept, sel.repeat(N)])
    new = regroup_rows(optimizer, src, n_keep)
    if n_new:
        scaling = new["scaling"].data[n_keep:]
        act = torch.exp(scaling.repeat(1, 3) if scaling.shape[1] == 1 else scaling)  # the parents' get_scaling
        z = samples.to(dev).float() if samples is not None else torch.randn((n_new, 3), device=dev, generator=generator)
        rots = build_rotation(new["rotation"].data[n_keep:])
        new["xyz"].data[n_keep:] = torch.bmm(rots, (z * act).unsqueeze(-1)).squeeze(-1) + new["xyz"].data[n_keep:]
        new_scaling = torch.log(act / (0.8 * N))
        new["scaling"].data[n_keep:] = new_scaling[:, :scaling.shape[1]]
    return new, _reset_stats(stats, src.numel(), dev)


@torch.no_grad()
def densify_and_prune(optimizer: FusedAdam, stats: Dict[str, torch.Tensor], max_grad, min_opacity, extent, max_screen_size,
                      percent_dense: float = 0.01, generator: Optional[torch.Generator] = None):
    """GaussianModel.densify_and_prune (:501-514) on ``stats`` = {xyz_gradient_accum (P,1), denom (P,1), max_radii2D (P)}."""
    grads = stats["xyz_gradient_accum"] / stats["denom"]
    grads[grads.isnan()] = 0.0
    _, stats = densify_and_clone(optimizer, stats, grads, max_grad, extent, percent_dense)
    _, stats = densify_and_split(optimizer, stats, grads, max_grad, extent, percent_dense, generator=generator)
    opacity = torch.sigmoid(_groups(optimizer)["opacity"]["params"][0].detach())
    prune_mask = (opacity < min_opacity).squeeze()
    if max_screen_size:
        big_points_vs = stats["max_radii2D"] > max_screen_size
        big_points_ws = _scaling_activated(optimizer).max(dim=1).values > 0.1 * extent
        prune_mask = torch.logical_or(torch.logical_or(prune_mask, big_points_vs), big_points_ws)
    return prune_points(optimizer, prune_mask, stats)
