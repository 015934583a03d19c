"""Stand-in for the reference's pybind module ``diff_gaussian_rasterization._C`` (dgr/ext.cpp:16-18).

Same three functions, same argument order and return tuples as ``RasterizeGaussiansCUDA``,
``RasterizeGaussiansBackwardCUDA`` and ``markVisible`` (dgr/rasterize_points.cu:35-229); the work is
done by libriggs_b200.so through its C ABI.  Tensors are only unwrapped to device pointers here.
"""
from __future__ import annotations

import torch

from .. import _lib
from .._lib import check, f32c, ptr, stream_ptr

NUM_CHANNELS = 3


def _on(device, t):
    """viewmatrix / campos etc. may arrive on another device in sloppy callers; the reference would fault."""
    if t is not None and t.numel() and not t.is_cuda:
        raise _lib.RiggsError("rasterize_gaussians: all non-empty tensors must live on the CUDA device")
    return t


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug):
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # AT_ERROR, rasterize_points.cu:57-59
    if not means3D.is_cuda:
        raise _lib.RiggsError("rasterize_gaussians needs CUDA tensors; riggs_b200 has no CPU path")
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    dev = means3D.device
    out_color = torch.empty((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
    out_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    out_alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    geom, binning, img = _lib.GrowBuffer(dev), _lib.GrowBuffer(dev), _lib.GrowBuffer(dev)
    rendered = 0
    if P == 0:  # rasterize_points.cu:83: nothing is launched, the images stay zero
        out_color.zero_(); out_depth.zero_(); out_alpha.zero_()
        return rendered, out_color, out_depth, out_alpha, radii, geom.t, binning.t, img.t
    M = sh.size(1) if sh.numel() != 0 else 0
    means3D, colors, opacity, scales, rotations, cov3D_precomp, sh = (f32c(t) for t in (
        means3D, colors, opacity, scales, rotations, cov3D_precomp, sh))
    background, viewmatrix, projmatrix, campos = (f32c(_on(dev, t)) for t in (background, viewmatrix, projmatrix, campos))
    with torch.cuda.device(dev):
        rendered = check(_lib.lib().rg_rasterize_forward(
            geom.fn, None, binning.fn, None, img.fn, None, P, int(degree), M, ptr(background), W, H, ptr(means3D),
            ptr(sh), ptr(colors), ptr(opacity), ptr(scales), float(scale_modifier), ptr(rotations),
            ptr(cov3D_precomp), ptr(viewmatrix), ptr(projmatrix), ptr(campos), float(tan_fovx), float(tan_fovy),
            int(bool(prefiltered)), ptr(out_color), ptr(out_depth), ptr(out_alpha), ptr(radii), int(bool(debug)),
            stream_ptr()), "rasterize_gaussians")
    return rendered, out_color, out_depth, out_alpha, radii, geom.t, binning.t, img.t


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                 dL_dout_depth, dL_dout_alpha, sh, degree, campos, geomBuffer, R, binningBuffer,
                                 imageBuffer, alphas, debug):
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    dev = means3D.device
    opts = dict(dtype=torch.float32, device=dev)
    # The kernels write every row, so no zero fill is needed (the reference zero-fills, rasterize_points.cu:158-167).
    dL_dmeans3D = torch.empty((P, 3), **opts)
    dL_dmeans2D = torch.empty((P, 3), **opts)
    dL_dcolors = torch.empty((P, NUM_CHANNELS), **opts)
    dL_ddepths = torch.empty((P, 1), **opts)
    dL_dconic = torch.empty((P, 2, 2), **opts)
    dL_dopacity = torch.empty((P, 1), **opts)
    dL_dcov3D = torch.empty((P, 6), **opts)
    dL_dsh = torch.empty((P, M, 3), **opts)
    dL_dscales = torch.empty((P, 3), **opts)
    dL_drotations = torch.empty((P, 4), **opts)
    if P != 0:
        have_scales = scales.numel() != 0
        if not have_scales:
            dL_dscales.zero_(); dL_drotations.zero_()
        means3D, colors, scales, rotations, cov3D_precomp, sh, alphas = (f32c(t) for t in (
            means3D, colors, scales, rotations, cov3D_precomp, sh, alphas))
        g_color, g_depth, g_alpha = f32c(dL_dout_color), f32c(dL_dout_depth), f32c(dL_dout_alpha)
        background, viewmatrix, projmatrix, campos = (f32c(t) for t in (background, viewmatrix, projmatrix, campos))
        radii = radii.contiguous()
        with torch.cuda.device(dev):
            check(_lib.lib().rg_rasterize_backward(
                P, int(degree), M, int(R), ptr(background), W, H, ptr(means3D), ptr(sh), ptr(colors), ptr(alphas),
                ptr(scales), float(scale_modifier), ptr(rotations), ptr(cov3D_precomp), ptr(viewmatrix),
                ptr(projmatrix), ptr(campos), float(tan_fovx), float(tan_fovy), ptr(radii), ptr(geomBuffer),
                ptr(binningBuffer), ptr(imageBuffer), ptr(g_color), ptr(g_depth), ptr(g_alpha), ptr(dL_dmeans2D),
                ptr(dL_dconic), ptr(dL_dopacity), ptr(dL_dcolors), ptr(dL_ddepths), ptr(dL_dmeans3D),
                ptr(dL_dcov3D), ptr(dL_dsh), ptr(dL_dscales) if have_scales else None,
                ptr(dL_drotations) if have_scales else None, int(bool(debug)), stream_ptr()),
                "rasterize_gaussians_backward")
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P != 0:
        if not means3D.is_cuda:
            raise _lib.RiggsError("mark_visible needs CUDA tensors; riggs_b200 has no CPU path")
        means3D, viewmatrix, projmatrix = f32c(means3D), f32c(viewmatrix), f32c(projmatrix)
        with torch.cuda.device(means3D.device):
            check(_lib.lib().rg_mark_visible(P, ptr(means3D), ptr(viewmatrix), ptr(projmatrix), present.data_ptr(),
                                             stream_ptr()), "mark_visible")
    return present


def state_ptrs(P, R, W, H, geomBuffer, binningBuffer, imageBuffer):
    """Test/debug helper: device addresses of the arrays inside the three opaque state buffers."""
    import ctypes as C
    out = (C.c_void_p * 13)()
    _lib.lib().rg_state_ptrs(P, int(R), W, H, ptr(geomBuffer), ptr(binningBuffer), ptr(imageBuffer), out)
    return [int(v) if v else 0 for v in out]


def debug_sorted_keys(P, R, W, H, geomBuffer, binningBuffer, imageBuffer):
    """Test helper: the reference's sorted (tile << 32 | depth bits) keys rebuilt from our state."""
    keys = torch.empty(int(R), dtype=torch.int64, device=geomBuffer.device)
    if R > 0:
        check(_lib.lib().rg_debug_sorted_keys(P, int(R), W, H, ptr(geomBuffer), ptr(binningBuffer), ptr(imageBuffer),
                                              ptr(keys), stream_ptr()), "debug_sorted_keys")
    return keys
