"""
How often does an edge of a discontinuous carrier under FM land one sample away from where the
serial fp64 recurrence (the oracle) puts it?  DESIGN.md 4.2 argues that under FM the position of
a discontinuity is decided by the rounding noise of upstream's own recurrence, so no other
evaluation order reproduces it to the sample; the tests allow a bounded number of such flips.
This measures the rate: voices of one carrier kind, FM by a Sine, rendered one by one through
the bank kernel and through the C oracle; a "flip" is a sample more than 1 LSB off (full scale).

  python tools/edge_flip_rate.py [voices per kind] [seconds]       (GPU box)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import synthesizer_b200 as sb                      # noqa: E402
from oracle import coracle as co                   # noqa: E402  (the checker, not the product)

nv = int(sys.argv[1]) if len(sys.argv) > 1 else 24
seconds = float(sys.argv[2]) if len(sys.argv) > 2 else 20.0
sr = 48000
n = int(sr * seconds)
rng = np.random.default_rng(2)
names = {2: "Square", 3: "Sawtooth", 4: "Pulse"}
classes = {2: sb.Square, 3: sb.Sawtooth, 4: sb.Pulse}
for cw in (2, 3, 4):
    flips = edges = 0
    for _ in range(nv):
        f0, fm, depth = float(rng.uniform(55, 3520)), float(rng.uniform(0.5, 220)), float(np.float32(rng.uniform(0.002, 0.05)))
        kw = dict(pulsewidth=0.3) if cw == 4 else {}
        car = classes[cw](f0, 0.99, 0.0, 0.0, fm_lfo=sb.Sine(fm, depth, samplerate=sr), samplerate=sr, **kw)
        got = sb.VoiceBank(sr).add(car, 1.0, 0.0).render(n).cpu().numpy()[0::2].astype(np.float64)
        lfo = co.osc(0, n, sr, fm, depth)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.0, 0.0, fm=lfo, pulsewidth=0.3)
        flips += int((np.abs(np.trunc(got * 32767) - np.trunc(want * 32767)) > 1).sum())
        edges += int((np.abs(np.diff(want)) > 0.5).sum())
    vs = nv * n
    print("%-8s carrier x Sine FM: %d flips in %.3g voice-samples (%d edges) = %.2f per 1e9 voice-samples, %.3g per edge"
          % (names[cw], flips, vs, edges, flips / vs * 1e9, flips / max(edges, 1)))
