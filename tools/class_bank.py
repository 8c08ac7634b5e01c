"""
Renders one 8192-voice bank of a single (carrier, modulator) kind a few times: a target for
`ncu -k regex:bank_` when looking at the edged or generic kernels (bench.py only exercises the
Sine x Sine kind).   python tools/class_bank.py <carrier 0..4> <modulator 0..4 | -1 | 5 | 6> [voices]
(modulator 5 = a Sine of FM gain 6 - 10 rad: the generic kernels; 6 = a Sine of gain 1.5 - 2.6 rad: the fast
path's strong-FM variant for Sine / Triangle carriers; carrier 9 = a uniform mix of all
5 x 5 (carrier, modulator) kinds, what BASELINE config 3 calls "mixed waveforms")
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthesizer_b200 as sb            # noqa: E402

cw, mw = int(sys.argv[1]), int(sys.argv[2])
nv = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
sr, nfr = 48000, 480000
classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
r = np.random.default_rng(11)
b = sb.VoiceBank(sr)
mixed = cw == 9
for v in range(nv):
    if mixed:
        cw, mw = v % 5, (v // 5) % 5
    f0 = float(r.uniform(55, 1760))
    if mw in (5, 6):
        lfo, f0 = sb.Sine(float(r.uniform(0.5, 12)), 16.0 if mw == 5 else 3.9, samplerate=sr), float(r.uniform(3000, 5000))
    else:
        lfo = classes[mw](float(r.uniform(0.5, 12)), float(r.uniform(0.005, 0.06)), samplerate=sr) if mw >= 0 else None
    car = classes[cw](f0, 0.9 / nv, fm_lfo=lfo, samplerate=sr)
    pan = float(r.uniform(-1, 1))
    b.add(sb.EnvelopeFilter(car, 0.05, 0.3, 8.0, 0.7, 1.5), (1 - pan) / 2, (1 + pan) / 2)
vd, _ = b._upload(0, None)
g, sm, e, sc = b.count_classes()
for _ in range(3):
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    b.render(nfr, voices_dev=vd, generic_voices=g, smooth_voices=sm, edged_voices=e, scan_voices=sc)
    t1.record()
    torch.cuda.synchronize()
    print("%.3f ms (generic %d, edged %d, scan %d)" % (t0.elapsed_time(t1), g, e, sc))
