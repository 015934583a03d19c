"""numpy/ctypes face of oracle/raster_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build_oracle

_libs = {}


def lib(f64: bool = False):
    """f64=False: liboracle.so (fp32, the restatement proper); f64=True: the same source built with -DORC_DOUBLE,
    every expression evaluated in double on double arrays (the arbiter for ill-conditioned gradients)."""
    L = _libs.get(bool(f64))
    if L is None:
        L = C.CDLL(build_oracle.build_cpu_oracle(f64=f64))
        L.orc_scan.restype = C.c_int64
        L.orc_higher_msb.restype = C.c_uint32
        L.orc_higher_msb.argtypes = [C.c_uint32]
        assert L.orc_real_bytes() == (8 if f64 else 4)
        _libs[bool(f64)] = L
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f(a, f64: bool = False):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64 if f64 else np.float32)


def _real(f64):
    return (np.float64, C.c_double) if f64 else (np.float32, C.c_float)


def forward(means3D, opacities, view, proj, campos, W, H, tanfovx, tanfovy, bg, shs=None, colors_precomp=None,
            scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0, degree=3, geom_override=None, f64=False):
    """Whole forward pipeline on the CPU.  Returns a dict with every intermediate the gates compare.

    geom_override: dict with radii/means2D/depths/tiles_touched/(conic_opacity, rgb) taken from another
    implementation (e.g. the GPU) -- the integer binning stages then run on exactly those floats."""
    L = lib(f64)
    ft, cf = _real(f64)
    _f = lambda a: globals()["_f"](a, f64)  # noqa: E731
    P = means3D.shape[0]
    means3D, opacities, view, proj, campos, bg = map(_f, (means3D, opacities, view, proj, campos, bg))
    shs, colors_precomp, scales, rotations, cov3D_precomp = map(_f, (shs, colors_precomp, scales, rotations, cov3D_precomp))
    M = shs.shape[1] if shs is not None else 0
    radii = np.zeros(P, np.int32)
    means2D = np.zeros((P, 2), ft)
    depths = np.zeros(P, ft)
    cov3D = np.zeros((P, 6), ft)
    rgb = np.zeros((P, 3), ft)
    conic_opacity = np.zeros((P, 4), ft)
    tiles = np.zeros(P, np.uint32)
    clamped = np.zeros((P, 3), np.uint8)
    L.orc_preprocess(P, degree, M, _p(means3D), _p(scales), cf(scale_modifier), _p(rotations), _p(opacities),
                     _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(view), _p(proj), _p(campos), W, H,
                     cf(tanfovx), cf(tanfovy), _p(radii), _p(means2D), _p(depths), _p(cov3D), _p(rgb),
                     _p(conic_opacity), _p(tiles), _p(clamped))
    if colors_precomp is not None:
        rgb = colors_precomp.copy()
    out = dict(radii=radii, means2D=means2D, depths=depths, cov3D=cov3D if cov3D_precomp is None else cov3D_precomp,
               rgb=rgb, conic_opacity=conic_opacity, tiles_touched=tiles, clamped=clamped)
    if geom_override is not None:
        for k, v in geom_override.items():
            out[k] = np.ascontiguousarray(v)
    out.update(binning(out["radii"], out["means2D"], out["depths"], out["tiles_touched"], W, H, f64=f64))
    out.update(blend_forward(out, W, H, bg, f64=f64))
    return out


def binning(radii, means2D, depths, tiles, W, H, f64=False):
    L = lib(f64)
    P = radii.shape[0]
    radii = np.ascontiguousarray(radii, np.int32)
    means2D, depths = _f(means2D, f64), _f(depths, f64)
    tiles = np.ascontiguousarray(tiles, np.uint32)
    offsets = np.zeros(P, np.uint32)
    R = int(L.orc_scan(P, _p(tiles), _p(offsets)))
    keys_u = np.zeros(R, np.uint64)
    vals_u = np.zeros(R, np.uint32)
    L.orc_duplicate_with_keys(P, W, H, _p(means2D), _p(depths), _p(offsets), _p(radii), _p(keys_u), _p(vals_u))
    gx, gy = (W + 15) // 16, (H + 15) // 16
    bit = int(L.orc_higher_msb(gx * gy))
    keys = np.zeros(R, np.uint64)
    vals = np.zeros(R, np.uint32)
    L.orc_sort_pairs(C.c_int64(R), _p(keys_u), _p(vals_u), _p(keys), _p(vals), 32 + bit)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    L.orc_tile_ranges(C.c_int64(R), _p(keys), gx * gy, _p(ranges))
    return dict(point_offsets=offsets, num_rendered=R, keys_unsorted=keys_u, vals_unsorted=vals_u, keys=keys,
                point_list=vals, ranges=ranges)


def blend_forward(st, W, H, bg, f64=False):
    L = lib(f64)
    ft, _ = _real(f64)
    _f = lambda a: globals()["_f"](a, f64)  # noqa: E731
    color = np.zeros((3, H, W), ft)
    depth = np.zeros((1, H, W), ft)
    alpha = np.zeros((1, H, W), ft)
    n_contrib = np.zeros((H, W), np.uint32)
    bg = _f(bg)
    L.orc_blend_forward(W, H, _p(st["ranges"]), _p(st["point_list"]), _p(_f(st["means2D"])), _p(_f(st["rgb"])),
                        _p(_f(st["depths"])), _p(_f(st["conic_opacity"])), _p(bg), _p(color), _p(depth), _p(alpha),
                        _p(n_contrib))
    return dict(color=color, depth=depth, alpha=alpha, n_contrib=n_contrib)


def backward(st, means3D, view, proj, campos, W, H, tanfovx, tanfovy, bg, dL_dcolor, dL_ddepth, dL_dalpha, shs=None,
             colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0, degree=3,
             f64=False):
    """Backward pipeline on the CPU from the forward state dict `st`."""
    L = lib(f64)
    ft, cf = _real(f64)
    _f = lambda a: globals()["_f"](a, f64)  # noqa: E731
    P = means3D.shape[0]
    means3D, view, proj, campos, bg = map(_f, (means3D, view, proj, campos, bg))
    shs, scales, rotations = map(_f, (shs, scales, rotations))
    M = shs.shape[1] if shs is not None else 0
    dmean2D = np.zeros((P, 2), np.float64)
    dconic = np.zeros((P, 3), np.float64)
    dopacity = np.zeros(P, np.float64)
    dcolor = np.zeros((P, 3), np.float64)
    ddepth = np.zeros(P, np.float64)
    gc, gd, ga = _f(dL_dcolor), _f(dL_ddepth), _f(dL_dalpha)
    colors = _f(st["rgb"])
    L.orc_blend_backward(P, W, H, _p(st["ranges"]), _p(st["point_list"]), _p(bg), _p(_f(st["means2D"])),
                         _p(_f(st["conic_opacity"])), _p(colors), _p(_f(st["depths"])), _p(_f(st["alpha"])),
                         _p(np.ascontiguousarray(st["n_contrib"], np.uint32)), _p(gc), _p(gd), _p(ga), _p(dmean2D),
                         _p(dconic), _p(dopacity), _p(dcolor), _p(ddepth))
    f_dmean2D, f_dconic, f_dcolor, f_ddepth = (x.astype(ft) for x in (dmean2D, dconic, dcolor, ddepth))
    dmean3D = np.zeros((P, 3), ft)
    dcov3D = np.zeros((P, 6), ft)
    dsh = np.zeros((P, M, 3), ft) if shs is not None else None
    dscale = np.zeros((P, 3), ft) if scales is not None else None
    drot = np.zeros((P, 4), ft) if scales is not None else None
    cov3D = _f(st["cov3D"])
    clamped = np.ascontiguousarray(st["clamped"], np.uint8)
    L.orc_preprocess_backward(P, degree, M, _p(means3D), _p(np.ascontiguousarray(st["radii"], np.int32)), _p(shs),
                              _p(clamped), _p(scales), _p(rotations), cf(scale_modifier), _p(cov3D), _p(view),
                              _p(proj), _p(campos), W, H, cf(tanfovx), cf(tanfovy), _p(f_dmean2D),
                              _p(f_dconic), _p(f_dcolor), _p(f_ddepth), _p(dmean3D), _p(dcov3D), _p(dsh), _p(dscale),
                              _p(drot))
    return dict(dmean2D=f_dmean2D, dconic=f_dconic, dopacity=dopacity.astype(ft), dcolor=f_dcolor,
                ddepth=f_ddepth, dmean3D=dmean3D, dcov3D=dcov3D, dsh=dsh, dscale=dscale, drot=drot)


def dist2_knn3(points):
    L = lib()
    pts = _f(points)
    out = np.zeros(pts.shape[0], np.float32)
    L.orc_dist2_knn3(pts.shape[0], _p(pts), _p(out))
    return out


def mark_visible(means3D, view, proj):
    L = lib()
    m = _f(means3D)
    out = np.zeros(m.shape[0], np.uint8)
    L.orc_mark_visible(m.shape[0], _p(m), _p(_f(view)), _p(_f(proj)), _p(out))
    return out.astype(bool)
