#!/usr/bin/env python
"""
Writes tests/golden/synthplayer_golden.npz from the REAL synthplayer package.

    SYNTHPLAYER_SRC=/path/to/checkout python tests/golden/make_synthplayer_golden.py

SYNTHPLAYER_SRC (default /root/reference) is a checkout of irmen/synthesizer (the directory that
holds ``synthplayer/``).  The script imports ``synthplayer.oscillators``, ``synthplayer.sample``
and ``synthplayer.synth`` from there, runs every case of tests/golden/synthplayer_cases.py with
them and stores the outputs plus a manifest (package version, git commit if any, interpreter,
the uniform draws white noise consumed, and — per case — the error if the case could not be built:
a wrong recalled signature shows up as a named failure, not as a crash).

After that, ``pytest tests/test_reference_golden.py`` compares the oracle (CPU) and, with
``-m gpu``, the CUDA path with the file; until the file exists those tests skip with the reason
"reference absent".  As of this commit /root/reference holds only README.md, so the file has
never been generated.

    --module oracle    dry run of the plumbing with oracle/ standing in for the package (used by
                       tests/test_reference_golden.py::test_generator_plumbing; never commit its output)
    --out PATH         where to write (default tests/golden/synthplayer_golden.npz)
"""
import argparse
import json
import os
import platform
import random
import subprocess
import sys
import traceback

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import synthplayer_cases as cases      # noqa: E402


class DrawRecorder:
    """Replaces random.uniform / random.random while a white-noise case renders, and records the
    uniform draws in [0, 1) that it consumed (random.uniform(a, b) is a + (b-a)*random())."""

    def __init__(self):
        self.draws = {}
        self._case = None
        self._rng = random.Random(20240229)
        self._orig = (random.uniform, random.random)

    def begin(self, case):
        self._case = case
        self.draws[case] = []

    def _random(self):
        u = self._rng.random()
        if self._case is not None:
            self.draws[self._case].append(u)
        return u

    def _uniform(self, a, b):
        return a + (b - a) * self._random()

    def install(self):
        random.uniform, random.random = self._uniform, self._random

    def remove(self):
        random.uniform, random.random = self._orig


def reference_namespace(src):
    if not os.path.isdir(os.path.join(src, "synthplayer")):
        raise SystemExit("no synthplayer/ under %s: the reference source is absent (set SYNTHPLAYER_SRC)" % src)
    rec = DrawRecorder()
    rec.install()                    # before the import, in case the package binds random.uniform by name
    sys.path.insert(0, src)
    import synthplayer
    import synthplayer.oscillators as osc
    import synthplayer.sample as smp
    import synthplayer.synth as syn
    ns = cases.Namespace("reference", osc, smp.Sample, syn.WaveSynth)
    ns.recorder = rec
    commit = None
    try:
        commit = subprocess.run(["git", "-C", src, "rev-parse", "HEAD"], capture_output=True, text=True,
                                timeout=10).stdout.strip() or None
    except Exception:
        pass
    info = {"package": "synthplayer", "version": getattr(synthplayer, "__version__", None), "commit": commit,
            "source": src}
    return ns, info


def oracle_namespace():
    from oracle import synth_oracle as so, sample_oracle as sa

    class _Rec:
        """the oracle takes injected draws; hand it seeded ones and record them like the real run"""
        def __init__(self):
            self.draws = {}

    draws = {}
    rng = random.Random(20240229)

    class NS(cases.Namespace):
        def noise(self, case, frequency, amplitude=1.0, bias=0.0, samplerate=cases.SR):
            draws[case] = [rng.random() for _ in range(4096)]
            return self.o.WhiteNoise(frequency, amplitude, bias, samplerate=samplerate, noise_source=draws[case])

    ns = NS("oracle", so, sa.Sample, sa.WaveSynth if hasattr(sa, "WaveSynth") else so.WaveSynth)
    rec = _Rec()
    rec.draws = draws
    ns.recorder = rec
    return ns, {"package": "oracle (dry run of the plumbing, NOT the reference)", "version": None, "commit": None,
                "source": "oracle/"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--module", default="synthplayer", choices=["synthplayer", "oracle"])
    ap.add_argument("--out", default=os.path.join(HERE, "synthplayer_golden.npz"))
    args = ap.parse_args()
    if args.module == "synthplayer":
        ns, info = reference_namespace(os.environ.get("SYNTHPLAYER_SRC", "/root/reference"))
    else:
        ns, info = oracle_namespace()
    arrays, manifest = {}, {"reference": info, "python": platform.python_version(), "cases": {}}
    for name in cases.OSC_CASES:
        try:
            arrays[name] = cases.render_osc(ns, name)
            manifest["cases"][name] = {"type": "osc", "frames": int(arrays[name].size)}
        except Exception as e:
            manifest["cases"][name] = {"type": "osc", "error": "%s: %s" % (type(e).__name__, e),
                                       "trace": traceback.format_exc(limit=3)}
    for name in cases.SAMPLE_CASES:
        try:
            vals, rate, width, nch = cases.render_sample(ns, name)
            arrays[name] = vals.astype(np.int32)
            manifest["cases"][name] = {"type": "sample", "samplerate": rate, "samplewidth": width, "nchannels": nch}
        except Exception as e:
            manifest["cases"][name] = {"type": "sample", "error": "%s: %s" % (type(e).__name__, e),
                                       "trace": traceback.format_exc(limit=3)}
    if hasattr(ns.recorder, "remove"):
        ns.recorder.remove()
    manifest["draws"] = {k: list(v) for k, v in ns.recorder.draws.items()}
    arrays["__manifest__"] = np.frombuffer(json.dumps(manifest).encode(), dtype=np.uint8)
    np.savez_compressed(args.out, **arrays)
    failed = [k for k, v in manifest["cases"].items() if "error" in v]
    print("wrote %s: %d cases, %d could not be built%s" % (args.out, len(manifest["cases"]), len(failed),
                                                            (": " + ", ".join(failed)) if failed else ""))


if __name__ == "__main__":
    main()
