"""BASELINE config 2 (FM sawtooth + ADSR + echo, 60 s stereo 48 kHz through the oscillator API),
a few times: a target for `ncu --metrics gpu__time_duration.sum` when looking at its launches."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthesizer_b200 as sb            # noqa: E402

sr = 48000


def config2():
    lfo = sb.Sine(5.0, 0.03, samplerate=sr)
    car = sb.Sawtooth(220.0, 0.45, fm_lfo=lfo, samplerate=sr)
    chain = sb.EchoFilter(sb.EnvelopeFilter(car, 0.5, 1.5, 40.0, 0.6, 10.0), 1.0, 6, 0.35, 0.5)
    return sb.Sample.from_oscillator(chain, 60.0).stereo()


for _ in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    s = config2()
    torch.cuda.synchronize()
    print("%.3f ms, %d frames" % ((time.perf_counter() - t0) * 1e3, len(s)))
