"""CPU-only: the reference arm of bench.py (`--impl reference`: the oracle port on the host cores) prints one JSON line
with the keys of the bench contract, on the same metric / unit / workload text as the library's own arm."""
import json
import os
import subprocess
import sys

import pytest

import helpers


@pytest.mark.timeout(300)
def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(helpers.ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=280, cwd=helpers.ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.strip().splitlines() if l.startswith("{")]
    assert len(lines) == 1  # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["higher_is_better"] is True and d["value"] > 0 and d["unit"] == "iters/s"
    assert abs(d["ms_per_step"] * d["value"] - 1000.0) < 1.0
    assert d["vs_baseline"] is None and d["data"] == "synthetic" and d["dtype"] == "f32"
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    # the same metric and workload text as the library's own arm prints for this configuration
    sys.path.insert(0, helpers.ROOT)
    import bench
    assert d["metric"] == bench.METRIC and d["unit"] == bench.UNIT
    assert d["config"]["workload"].startswith("C0: 300000 Gaussians, 24-joint skeleton, 800x800, 1 view/GPU/step")
