"""bench.py's output contract, checked on the GPU box: one JSON line on stdout with the keys
the driver reads, numbers in plausible ranges, and a non-zero count of this library's kernel
launches in the timed region (a short run: 2 steps, no CPU baseline, no extra legs)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_prints_one_json_line_with_the_contract_keys(cuda):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup", "1",
                          "--no-cpu-baseline", "--no-extras"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, "stdout must carry the JSON line and nothing else: %r" % lines[:3]
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "cpu_baseline"):
        assert key in d, key
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] >= 3 and d["higher_is_better"] is True
    assert d["scaling"] == "strong" and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["config"]["voices_total"] == 65536                 # BASELINE config 5 exactly, on one GPU
    assert d["value"] > 1e5 and 0.5 < d["ms_per_step"] < 100
    assert d["weak"]["voices_per_gpu"] == 8192 and 0.5 < d["weak"]["ms_per_step"] < 20
    assert d["gpu_launches"] >= 2 * 3, "the timed steps must launch this library's kernels"
    e = d["e2e"]
    assert e["unit"] == d["unit"] and 0 < e["value"] <= d["value"] * 1.02
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    r = d["roofline"]
    assert r["bound"] in ("sfu", "hbm", "tensor") and r["peak"] > 0 and 0 < r["frac"] < 1.2
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    # the dominant kernel is timed alone, inside the timed region, and fits inside the step
    assert 0 < r["kernel_ms"] < d["ms_per_step"] and 0.5 < r["kernel_share_of_step"] < 1.0
    assert 0 < r["hardware"]["mufu_frac"] <= r["frac"]
    c = d["clocks"]
    assert set(("sm_mhz", "sm_max_mhz", "reasons")) <= set(c)
