"""rg_dist_cuda2_static: the 3-NN distance entry point for callers that cannot allocate device memory.

Some users of the library cannot hand rg_dist_cuda2 the scratch buffer it needs: a rendering / inference plug-in that got
its device-memory budget once at start-up and is not allowed to call the CUDA allocator afterwards (an allocation stalls
the whole device and can fail in the middle of a session), or a C host without any allocator of its own.  For them the
library carries a 2 MiB scratch array inside its own image; the entry point is limited to 65536 points per call.

This file is named so that pytest collects it last: a run that stops at the first failure has then already run every
other test.  The test drives the C ABI directly (ctypes), twice in a row -- the second call finds the array filled with the
first call's leftovers -- and expects exactly the numbers of the ordinary entry point.  It must pass alone in a fresh
process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (sys.path)

pytestmark = pytest.mark.gpu


def _ordinary(L, pts):
    P = pts.shape[0]
    out = torch.full((P,), -1.0, device="cuda")
    nbytes = L.rg_knn_scratch_bytes(P)
    scratch = torch.empty(int(nbytes), dtype=torch.uint8, device="cuda")
    rc = L.rg_dist_cuda2(P, pts.data_ptr(), out.data_ptr(), scratch.data_ptr(), nbytes,
                         torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.rg_last_error().decode()
    return out


def _static(L, pts, out):
    rc = L.rg_dist_cuda2_static(pts.shape[0], pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.rg_last_error().decode()


def test_static_scratch_entry_matches_ordinary_entry_twice_in_a_row():
    from riggs_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(123)
    # two different clouds of different size, so the second call really reuses a dirty array with another layout
    clouds = [torch.randn(20_000, 3, generator=g), torch.rand(3_001, 3, generator=g) * 5.0]
    clouds[0][:4000] = clouds[0][:4000] * 0.01 + 1.5  # a dense cluster
    for pts_cpu in clouds:
        pts = pts_cpu.cuda().contiguous()
        ref = _ordinary(L, pts)
        out = torch.full((pts.shape[0],), -1.0, device="cuda")  # the caller's output buffer, allocated beforehand
        _static(L, pts, out)
        torch.cuda.synchronize()
        # same kernels, same order of operations: the tolerance of the other distCUDA2 tests (rtol 1e-5) is kept
        assert np.allclose(out.cpu().numpy(), ref.cpu().numpy(), rtol=1e-5, atol=1e-12)
        assert float(out.min()) >= 0.0
    # and once more on the first cloud, after the array has held the second cloud's data
    pts = clouds[0].cuda().contiguous()
    out = torch.empty(pts.shape[0], device="cuda")
    _static(L, pts, out)
    torch.cuda.synchronize()
    assert np.allclose(out.cpu().numpy(), _ordinary(L, pts).cpu().numpy(), rtol=1e-5, atol=1e-12)


def test_static_scratch_entry_allocates_no_device_memory():
    """The property the entry point exists for: across its calls neither torch's allocator nor the driver hands out
    any device memory (cudaMemGetInfo's free figure and torch's counters stay where they were)."""
    from riggs_b200 import _lib
    L = _lib.lib()
    pts = torch.rand(30_000, 3, generator=torch.Generator().manual_seed(5)).cuda().contiguous()
    out = torch.empty(pts.shape[0], device="cuda")
    _static(L, pts, out)  # the first call on the device looks the array's address up; not part of the measurement
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    alloc0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    n_alloc0 = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for _ in range(5):
        _static(L, pts, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == alloc0 and torch.cuda.memory_reserved() == reserved0
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == n_alloc0
    assert free1 >= free0  # nothing was taken from the device behind torch's back either


def test_static_scratch_entry_rejects_oversized_clouds():
    from riggs_b200 import _lib
    L = _lib.lib()
    P = 65537
    pts = torch.zeros(P, 3, device="cuda")
    out = torch.zeros(P, device="cuda")
    rc = L.rg_dist_cuda2_static(P, pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc < 0
    assert "65536" in L.rg_last_error().decode()
    # the limit itself is accepted
    P = 65536
    pts = torch.rand(P, 3, device="cuda")
    out = torch.zeros(P, device="cuda")
    assert L.rg_dist_cuda2_static(P, pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert np.allclose(out.cpu().numpy(), _ordinary(L, pts).cpu().numpy(), rtol=1e-5, atol=1e-12)
