"""CPU-only: the C-ABI library loads and exports every symbol include/riggs_b200.h declares."""
import ctypes
import os
import re

import helpers  # noqa: F401  (sys.path)
from riggs_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "riggs_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rg_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported():
    names = _declared()
    assert len(names) >= 14
    L = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/riggs_b200.h but not exported"


def test_binding_covers_header():
    assert sorted(_lib.exported_symbols()) == _declared()


def test_size_queries_without_gpu():
    L = _lib.lib()
    assert L.rg_version() >= 1
    g1, g2 = L.rg_geom_bytes(1000, 800, 800), L.rg_geom_bytes(2000, 800, 800)
    assert 0 < g1 < g2
    assert L.rg_image_bytes(800, 800) >= 800 * 800 * 4 + 2500 * 8
    assert L.rg_binning_bytes(10**6) >= 10**6 * 4
    assert L.rg_binning_bytes(0) > 0
    assert L.rg_knn_scratch_bytes(100000) >= 100000 * 16


def test_no_cpu_fallback():
    import pytest
    import torch
    from riggs_b200.diff_gaussian_rasterization import _C
    from riggs_b200.simple_knn._C import distCUDA2
    with pytest.raises(_lib.RiggsError):
        distCUDA2(torch.zeros(4, 3))
    e = torch.empty(0)
    with pytest.raises(_lib.RiggsError):
        _C.rasterize_gaussians(torch.zeros(3), torch.zeros(4, 3), e, torch.ones(4, 1), torch.ones(4, 3),
                               torch.ones(4, 4), 1.0, e, torch.eye(4), torch.eye(4), 0.5, 0.5, 16, 16,
                               torch.zeros(4, 16, 3), 3, torch.zeros(3), False, False)


def test_scene_struct_layout_matches_the_header():
    """RgSkinnedScene in riggs_b200/_lib.py mirrors the header field by field, in order (fields are only ever appended)."""
    import re
    text = open(os.path.join(ROOT, "include", "riggs_b200.h")).read()
    body = text[text.index("typedef struct RgSkinnedScene {"):text.index("} RgSkinnedScene;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.replace("typedef struct RgSkinnedScene {", "").strip()
        if not decl:
            continue
        parts = [p.strip() for p in decl.split(",")]
        names.append(re.split(r"[\s\*]+", parts[0])[-1])
        names += [p.lstrip("* ").strip() for p in parts[1:]]
    assert names == [n for n, _ in _lib.RgSkinnedScene._fields_]


def test_render_result_computes_the_visibility_filter_on_demand():
    import pytest
    import torch
    from riggs_b200.fused import _RenderResult
    out = _RenderResult({"radii": torch.tensor([0, 3, 0, 7], dtype=torch.int32), "render": torch.zeros(3, 2, 2)})
    assert "visibility_filter" in out and not dict.__contains__(out, "visibility_filter")
    vis = out["visibility_filter"]
    assert vis.tolist() == [False, True, False, True] and out.get("visibility_filter") is vis
    assert out.get("missing", 5) == 5 and "missing" not in out
    with pytest.raises(KeyError):
        out["missing"]
