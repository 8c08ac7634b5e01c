"""
Worst error of the bank's fast path against the C oracle as the FM gain grows (carrier radians per
unit of modulator output, 2*pi*f0*depth/sr): one FULL-SCALE voice per (carrier, gain, f0, modulator
frequency), 2 s at 48 kHz, error in int16 LSB.  The limits kFastMaxFmGain / kFastHighFmGain in
sb_voicebank.cu come from this table (a library built with -DFAST_MAX_FM_GAIN=8.5 shows the plain
run beyond its limit).   python tools/fm_gain_sweep.py
"""
import itertools
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthesizer_b200 as sb            # noqa: E402
from oracle import coracle as co         # noqa: E402

sr, n = 48000, 96000
print("carrier gain_rad f0 fm depth class max_lsb")
for cw in (0, 1):
    for gain in (1.0, 1.5, 2.0, 3.0, 4.0, 6.0, 8.0):
        for f0, fm in itertools.product((3517.9, 9973.1), (5.3, 109.7, 1999.3)):
            d = gain * sr / (2 * np.pi * f0)
            lfo = sb.Sine(fm, d, 0.3, samplerate=sr)
            car = [sb.Sine, sb.Triangle][cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
            bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
            kind = "generic" if bank.count_generic() else "fast"
            got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
            m = co.osc(0, n, sr, fm, d, 0.3, 0.0)
            want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
            err = np.abs(got - want) * 32767.0
            print("%s %.1f %.1f %.1f %.2f %s %.4f" % (["Sine", "Triangle"][cw], gain, f0, fm, d, kind, err.max()))
