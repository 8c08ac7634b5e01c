"""bench.py --impl reference runs without a GPU: it times the CPU port of the path (the
reference itself is absent offline) on the host cores and prints the same JSON line, marked
"impl": "reference".  One short step here."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_its_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines[:3]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["value"] > 0
    assert d["metric"].startswith("Msamples/s") and d["unit"] == "Msamples/s"
    assert "workload" in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_drives_the_real_package_when_it_is_there(tmp_path):
    """With a synthplayer package under baseline/_ref or SYNTHPLAYER_SRC the arm times THAT and
    says kind "reference".  The package is absent offline, so a stand-in with the same module
    layout (re-exporting the oracle's classes) plays its part here."""
    pkg = tmp_path / "synthplayer"
    pkg.mkdir()
    (pkg / "__init__.py").write_text("")
    (pkg / "oscillators.py").write_text(
        "import sys\nsys.path.insert(0, %r)\nfrom oracle.synth_oracle import *  # noqa\n" % ROOT)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=900, cwd=ROOT,
                         env={**os.environ, "SYNTHPLAYER_SRC": str(tmp_path)})
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["cpu_baseline"]["kind"] == "reference" and d["value"] > 0
    assert "baseline/_ref" in d["note"]
