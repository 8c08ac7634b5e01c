#!/usr/bin/env python
"""
Accuracy sweep of the voice-bank fast path against the oracle (test infrastructure; run on a
GPU box): one full-scale voice per (carrier f0, modulator f, FM depth), 2 s at 48 kHz, worst
error in int16 LSB.  Prints a table; the limits in voice_is_fast() and DESIGN.md come from it.
"""
import itertools
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthesizer_b200 as sb            # noqa: E402
from oracle import coracle as co         # noqa: E402

sr, n = 48000, 96000
rows = []
for cw in (0, 1, 3):
    for f0, fm, d in itertools.product((55.1, 441.3, 3517.9, 9973.0), (0.1, 5.3, 109.7, 1999.3), (0.01, 0.1, 0.5, 2.0)):
        lfo = sb.Sine(fm, d, 0.3, samplerate=sr)
        car = [sb.Sine, sb.Triangle, None, sb.Sawtooth][cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        fast = bank.count_generic() == 0
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, fm, d, 0.3, 0.0)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        # a flipped sawtooth edge is a full-scale error on one sample: report edges separately
        big = int((err > 100).sum())
        rows.append((cw, f0, fm, d, fast, float(np.percentile(err, 99.9)), float(err[err <= 100].max()), big))
print("wave f0 fm depth fast p99.9_lsb max_lsb(excluding edge flips) edge_flips")
for r in rows:
    print("%d %8.1f %7.1f %5.2f %d %8.4f %8.4f %d" % r)
worst = max(r[6] for r in rows if r[4])
print("worst non-edge error on the fast path: %.4f LSB" % worst)
per_rad = max(r[6] / (r[1] * 2 * np.pi / sr * r[3]) for r in rows if r[4] and r[1] * 2 * np.pi / sr * r[3] > 0.2)
print("worst error per radian of FM gain (fast voices with gain > 0.2 rad): %.3f LSB/rad" % per_rad)
print("worst non-edge error on the generic path: %.4f LSB" % max([r[6] for r in rows if not r[4]] or [0.0]))
