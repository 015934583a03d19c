"""ctypes face of oracle/_ref/libref_riggs.so, the reference's OWN CUDA kernels compiled unmodified
(see oracle/build_oracle.py and oracle/ref_shim.cu) -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Needs a GPU.  Used by the `-m gpu` tests as the primary oracle (bit-exact keys / order / ranges, image
and gradient tolerances) and by bench.py for the "reference CUDA on one B200" context number.
The reference launches on the legacy default stream: call it from torch's default stream only.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libref_riggs.so")
_lib = None
_p, _i, _f = C.c_void_p, C.c_int, C.c_float


def available() -> bool:
    return os.path.exists(LIB) and torch.cuda.is_available()


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB)
        L.ref_state_new.restype = _p
        L.ref_state_free.argtypes = [_p]
        L.ref_forward.restype = _i
        L.ref_forward.argtypes = [_p, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _i,
                                  _p, _p, _p, _p, _i]
        L.ref_backward.restype = _i
        L.ref_backward.argtypes = [_p, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f,
                                   _p] + [_p] * 13 + [_i]
        L.ref_state_ptrs.argtypes = [_p, C.POINTER(_p)]
        L.ref_mark_visible.argtypes = [_i, _p, _p, _p, _p]
        L.ref_dist_cuda2.argtypes = [_i, _p, _p]
        _lib = L
    return _lib


def _ptr(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _d2h(dev_ptr, nbytes):
    buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    C.CDLL("libcudart.so.12").cudaMemcpy(C.c_void_p(buf.data_ptr()), C.c_void_p(dev_ptr), C.c_size_t(nbytes), 3)
    return buf.cpu().numpy()


class RefRasterizer:
    def __init__(self):
        self.state = lib().ref_state_new()

    def __del__(self):
        try:
            lib().ref_state_free(self.state)
        except Exception:
            pass

    def forward(self, bg, means3D, opacities, view, proj, campos, W, H, tanfovx, tanfovy, shs=None,
                colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0, degree=3):
        P = means3D.shape[0]
        M = shs.shape[1] if shs is not None else 0
        dev = means3D.device
        color = torch.zeros((3, H, W), device=dev)
        depth = torch.zeros((1, H, W), device=dev)
        alpha = torch.zeros((1, H, W), device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        R = lib().ref_forward(self.state, P, degree, M, _ptr(bg), W, H, _ptr(means3D), _ptr(shs), _ptr(colors_precomp),
                              _ptr(opacities), _ptr(scales), scale_modifier, _ptr(rotations), _ptr(cov3D_precomp),
                              _ptr(view), _ptr(proj), _ptr(campos), tanfovx, tanfovy, 0, _ptr(color), _ptr(depth),
                              _ptr(alpha), _ptr(radii), 0)
        torch.cuda.synchronize()
        if R < 0:
            raise RuntimeError("reference forward failed")
        self.P, self.R, self.W, self.H = P, R, W, H
        return dict(num_rendered=R, color=color, depth=depth, alpha=alpha, radii=radii)

    def state_arrays(self):
        """Decode the reference's opaque buffers into numpy arrays (layouts of rasterizer_impl.cu:155-193)."""
        out = (_p * 16)()
        lib().ref_state_ptrs(self.state, out)
        P, R, HW = self.P, self.R, self.W * self.H
        T = ((self.W + 15) // 16) * ((self.H + 15) // 16)
        g = lambda i, n, dt: _d2h(out[i], n).view(dt) if out[i] else np.zeros(0, dt)
        d = dict(depths=g(0, 4 * P, np.float32), clamped=g(1, 3 * P, np.uint8).reshape(P, 3),
                 means2D=g(3, 8 * P, np.float32).reshape(P, 2), cov3D=g(4, 24 * P, np.float32).reshape(P, 6),
                 conic_opacity=g(5, 16 * P, np.float32).reshape(P, 4), rgb=g(6, 12 * P, np.float32).reshape(P, 3),
                 tiles_touched=g(7, 4 * P, np.uint32), point_offsets=g(8, 4 * P, np.uint32),
                 n_contrib=g(14, 4 * HW, np.uint32).reshape(self.H, self.W), ranges=g(15, 8 * T, np.uint32).reshape(T, 2))
        if R > 0:
            d.update(point_list=g(10, 4 * R, np.uint32), vals_unsorted=g(11, 4 * R, np.uint32),
                     keys=g(12, 8 * R, np.uint64), keys_unsorted=g(13, 8 * R, np.uint64))
        return d

    def backward(self, bg, means3D, radii, alpha, view, proj, campos, tanfovx, tanfovy, g_color, g_depth, g_alpha,
                 shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0,
                 degree=3):
        P, W, H = self.P, self.W, self.H
        M = shs.shape[1] if shs is not None else 0
        dev = means3D.device
        z = lambda *s: torch.zeros(s, device=dev)
        out = dict(dmean2D=z(P, 3), dconic=z(P, 4), dopacity=z(P, 1), dcolor=z(P, 3), ddepth=z(P, 1), dmean3D=z(P, 3),
                   dcov3D=z(P, 6), dsh=z(P, M, 3), dscale=z(P, 3), drot=z(P, 4))
        torch.cuda.synchronize()
        rc = lib().ref_backward(self.state, P, degree, M, _ptr(bg), W, H, _ptr(means3D), _ptr(shs), _ptr(colors_precomp),
                                _ptr(alpha), _ptr(scales), scale_modifier, _ptr(rotations), _ptr(cov3D_precomp),
                                _ptr(view), _ptr(proj), _ptr(campos), tanfovx, tanfovy, _ptr(radii), _ptr(g_color),
                                _ptr(g_depth), _ptr(g_alpha), _ptr(out["dmean2D"]), _ptr(out["dconic"]),
                                _ptr(out["dopacity"]), _ptr(out["dcolor"]), _ptr(out["ddepth"]), _ptr(out["dmean3D"]),
                                _ptr(out["dcov3D"]), _ptr(out["dsh"]) if M else None, _ptr(out["dscale"]),
                                _ptr(out["drot"]), 0)
        torch.cuda.synchronize()
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return out


def mark_visible(means3D, view, proj):
    P = means3D.shape[0]
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    torch.cuda.synchronize()
    lib().ref_mark_visible(P, _ptr(means3D), _ptr(view), _ptr(proj), present.data_ptr())
    torch.cuda.synchronize()
    return present


def dist_cuda2(points):
    P = points.shape[0]
    out = torch.zeros(P, device=points.device)
    torch.cuda.synchronize()
    lib().ref_dist_cuda2(P, _ptr(points), _ptr(out))
    torch.cuda.synchronize()
    return out
