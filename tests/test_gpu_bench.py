"""bench.py contract: ONE JSON line on stdout with the keys the driver reads, on a small workload."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

REQUIRED = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "roofline", "clocks", "gpu_launches", "e2e")


def _run(*extra):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup", "3", *extra],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [x for x in r.stdout.splitlines() if x.strip()]
    assert len(lines) == 1, r.stdout[-2000:]
    return json.loads(lines[0])


def test_bench_line_small_2d():
    d = _run("--workload", "c1", "--no-cpu-baseline")
    for k in REQUIRED:
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["value"] > 0
    assert d["config"]["workload"].startswith("c1") and d["data"] == "synthetic" and d["vs_baseline"] is None
    roof = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "kernel"):
        assert k in roof, k
    assert roof["bound"] == "hbm" and abs(roof["frac"] - roof["achieved"] / roof["peak"]) < 1e-12
    assert {r["kernel"] for r in d["roofline_kernels"]} == {"voxel_pass_fast", "rf_score"}
    assert roof["launch_ms"] == max(r["launch_ms"] for r in d["roofline_kernels"])
    assert d["gpu_launches"] > 0
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 8 * 512 * 512 and e["d2h_bytes_per_step"] > 0
    assert set(d["stage_ms"]) == {"table_init", "voxel_pass", "graph", "featurize", "rf_score"}


def test_bench_reference_arm_line():
    d = _run("--workload", "c1", "--impl", "reference")
    assert d["impl"] == "reference" and d["value"] > 0 and d["unit"] == "Mvoxels/s"
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
