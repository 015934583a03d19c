"""ctypes binding of libriggs_b200.so (the C ABI declared in include/riggs_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at first use, and every
wrapper raises when handed a non-CUDA tensor.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RG_LIB") or os.path.join(_HERE, "lib", "libriggs_b200.so")  # RG_LIB: a tuning build (build.py --variant)

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)
_p = C.c_void_p
_f = C.c_float
_i = C.c_int



class RgSkinnedScene(C.Structure):
    _fields_ = ([(n, C.c_int) for n in ("P", "J", "sh_degree", "M_rest", "isotropic", "zero_d_rotation", "width", "height")] +
                [(n, C.c_float) for n in ("tan_fovx", "tan_fovy", "scale_modifier")] +
                [(n, C.c_void_p) for n in ("xyz", "scaling_raw", "rotation_raw", "opacity_raw", "features_dc",
                                           "features_rest", "motion_mask", "mlp_weight", "offsets", "joints", "parents",
                                           "node_radius_raw", "local_rot", "global_trans", "viewmatrix", "projmatrix",
                                           "campos", "background")] +
                [("blend_mode", C.c_int), ("split_index", C.c_int), ("split_count", C.c_int)])


class RgAdamGroup(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_size_t), ("lr", C.c_float)]


class RgRowTensor(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_size_t)]


class RgWorkspace(C.Structure):
    _fields_ = [("geom", C.c_void_p), ("geom_bytes", C.c_size_t), ("image", C.c_void_p), ("image_bytes", C.c_size_t),
                ("binning", C.c_void_p), ("binning_bytes", C.c_size_t)]


_SIGNATURES = {
    "rg_version": (C.c_int, []),
    "rg_last_error": (C.c_char_p, []),
    "rg_geom_bytes": (C.c_size_t, [_i, _i, _i]),
    "rg_image_bytes": (C.c_size_t, [_i, _i]),
    "rg_binning_bytes": (C.c_size_t, [C.c_int64]),
    "rg_rasterize_forward": (C.c_int, [ALLOC_FN, _p, ALLOC_FN, _p, ALLOC_FN, _p, _i, _i, _i, _p, _i, _i, _p, _p, _p,
                                       _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _i, _p, _p, _p, _p, _i, _p]),
    "rg_rasterize_backward": (C.c_int, [_i, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f,
                                        _f, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i,
                                        _p]),
    "rg_mark_visible": (C.c_int, [_i, _p, _p, _p, _p, _p]),
    "rg_state_ptrs": (C.c_int, [_i, C.c_int64, _i, _i, _p, _p, _p, C.POINTER(C.c_void_p)]),
    "rg_debug_sorted_keys": (C.c_int, [_i, C.c_int64, _i, _i, _p, _p, _p, _p, _p]),
    "rg_debug_depth_sort_bytes": (C.c_size_t, [_i]),
    "rg_debug_depth_sort": (C.c_int, [_i, _p, _p, _p, _p, _p]),
    "rg_knn_scratch_bytes": (C.c_size_t, [_i]),
    "rg_dist_cuda2": (C.c_int, [_i, _p, _p, _p, C.c_size_t, _p]),
    "rg_dist_cuda2_static": (C.c_int, [_i, _p, _p, _p]),
    "rg_fk_forward": (C.c_int, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    "rg_fk_backward": (C.c_int, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "rg_binning_capacity": (C.c_uint32, [C.c_size_t]),
    "rg_skinned_forward": (C.c_int, [_p, _p] + [_p] * 9),
    "rg_skinned_status": (C.c_int, [_p, _p, _p, _p]),
    "rg_skinned_backward": (C.c_int, [_p, _p] + [_p] * 20),
    "rg_skinned_backward_part": (C.c_int, [_p, _p, _i, _i, _i] + [_p] * 20),
    "rg_skin_forward": (C.c_int, [_i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _p]),
    "rg_skin_backward": (C.c_int, [_i, _i] + [_p] * 19),
    "rg_skin_backward_blend": (C.c_int, [_i, _i] + [_p] * 10 + [_i] + [_p] * 9),
    "rg_l1_loss_scratch_bytes": (C.c_size_t, []),
    "rg_l1_loss": (C.c_int, [C.c_size_t, _p, _p, _p, _p, _p, C.c_size_t, _p]),
    "rg_allreduce_multimem": (C.c_int, [_p, C.c_size_t, _i, _i, _p]),
    "rg_allreduce_p2p": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t, _i, _i, _p]),
    "rg_allreduce_ranges": (C.c_int, [C.POINTER(C.c_void_p), _p, _i, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), _i, _i, _i, _p]),
    "rg_ssim_loss_scratch_bytes": (C.c_size_t, [_i, _i, _i]),
    "rg_ssim_l1_loss": (C.c_int, [_i, _i, _i, _p, _p, _f, _p, _p, _p, C.c_size_t, _p]),
    "rg_adam_step": (C.c_int, [_i, _p, _i, C.c_double, C.c_double, C.c_double, _p]),
    "rg_densification_stats": (C.c_int, [_i, _p, _p, _p, _p, _p, _p]),
    "rg_compact_scratch_bytes": (C.c_size_t, [_i]),
    "rg_compact_rows": (C.c_int, [_i, _p, _i, _p, _p, C.c_size_t, _p, _p]),
    "rg_gather_rows": (C.c_int, [C.c_longlong, _p, _i, _p, _p]),
    "rg_knn_points": (C.c_int, [_i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "rg_skin_topk_forward": (C.c_int, [_i, _i, _i] + [_p] * 15),
    "rg_skin_topk_backward": (C.c_int, [_i, _i, _i] + [_p] * 18),
    "rg_node_warp_forward": (C.c_int, [_i, _i, _i, _i] + [_p] * 12 + [_i] + [_p] * 9),
    "rg_node_warp_grad_channels": (C.c_int, []),
    "rg_node_warp_backward": (C.c_int, [_i, _i, _i, _i] + [_p] * 23),
    "rg_mlp_packed_bytes": (C.c_size_t, [_i]),
    "rg_mlp_pack": (C.c_int, [_i, _i, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), _p, _p]),
    "rg_mlp_forward": (C.c_int, [_i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, C.c_longlong, _p]),
    "rg_mlp_backward": (C.c_int, [_i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, C.c_longlong, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                  C.POINTER(C.c_int), C.POINTER(C.c_int), _p, _p]),
    "rg_launch_count": (C.c_longlong, []),
    "rg_profile_enable": (None, [_i]),
    "rg_profile_reset": (None, []),
    "rg_profile_count": (C.c_int, []),
    "rg_debug_tile_profile": (C.c_int, [_p, _p]),
    "rg_profile_get": (C.c_int, [_i, C.c_char_p, _i, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
}

_lib: Optional[C.CDLL] = None


def exported_symbols():
    """Names of every entry point include/riggs_b200.h declares."""
    return list(_SIGNATURES)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m riggs_b200.build` "
                "(riggs_b200 has no CPU or PyTorch fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class RiggsError(RuntimeError):
    pass


def check(rc: int, what: str) -> int:
    if rc < 0:
        raise RiggsError(f"{what}: {lib().rg_last_error().decode()}")
    return rc


def ptr(t: Optional[torch.Tensor]):
    """Device address of a tensor, or NULL for None / empty tensors (the reference's null-pointer convention)."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RiggsError("riggs_b200 kernels need CUDA tensors; there is no CPU path")
    return t.data_ptr()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def f32c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """contiguous fp32 view, 16-byte aligned (clones when a view is mis-aligned)."""
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.numel() and t.data_ptr() % 16:
        t = t.clone()
    return t


def profile_read():
    """{kernel name: (total_ms, launches)} of everything recorded since the last reset."""
    L = lib()
    out = {}
    for i in range(L.rg_profile_count()):
        name = C.create_string_buffer(64)
        ms, n = C.c_float(), C.c_int()
        if L.rg_profile_get(i, name, 64, C.byref(ms), C.byref(n)) == 0:
            out[name.value.decode()] = (ms.value, n.value)
    return out


class GrowBuffer:
    """A growable uint8 CUDA tensor handed to the library as an rg_alloc_fn, the analogue of the reference's
    resizeFunctional lambdas (dgr/rasterize_points.cu:27-33)."""

    def __init__(self, device):
        self.t = torch.empty(0, dtype=torch.uint8, device=device)

        def _alloc(_ctx, nbytes):
            if self.t.numel() < nbytes:
                self.t = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
            return self.t.data_ptr()

        self.fn = ALLOC_FN(_alloc)
