"""
Randomised parity soak of the voice-bank kernels against the C oracle (GPU box): banks of random size,
kinds, envelopes and start offsets, every (carrier, modulator) kind mixed, some voices forced to the
generic kernels, int16 PCM within 1 LSB except for the bounded share of edge flips the tests allow.
  python tools/bank_soak.py [rounds=24] [seed=1]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import test_gpu_voicebank as T            # noqa: E402  (make_bank, force_some_generic, oracle_bank_pcm)
from helpers import dev_pcm16             # noqa: E402

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 24
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
worst = 0.0
for r in range(rounds):
    nv = int(rng.integers(65, 700))
    secs = float(rng.uniform(0.05, 0.6))
    sr = int(rng.choice([44100, 48000]))
    start = int(rng.choice([0, 0, 1, 511, 512, 40000, 123457]))
    n = int(sr * secs) + int(rng.integers(0, 600))
    waves = None if rng.random() < 0.6 else list(rng.choice(5, size=int(rng.integers(1, 4)), replace=False))
    bank, P = T.make_bank(nv, sr, int(rng.integers(1, 1 << 30)), max(secs, 0.2) + start / sr, waves=waves, fm=True)
    for v in bank._voices:
        if v.m_wave == 4:
            v.m_pw = 0.1
    if rng.random() < 0.5:
        T.force_some_generic(bank, P)
    # a few strong-FM voices
    for i, v in enumerate(bank._voices):
        if v.m_wave >= 0 and v.c_wave < 2 and rng.random() < 0.1:
            v.m_amp = P[i, 8] = float(np.float32(rng.uniform(1.6, 5.5) * sr / (2 * np.pi * v.c_freq)))
    bank._dev = None
    got = dev_pcm16(bank.render(n, start=start))
    left, right = __import__("oracle.coracle", fromlist=["x"]).voice_bank(P, start + n, sr)
    want = np.empty(2 * n, dtype=np.int64)
    want[0::2] = T.pcm16_of(left[start:])
    want[1::2] = T.pcm16_of(right[start:])
    bad = np.abs(got.astype(np.int64) - want) > 1
    g, sm, e, sc = bank.count_classes()
    worst = max(worst, float(bad.mean()))
    print("round %2d: %3d voices (generic %d smooth %d edged %d scan %d) sr %d start %6d frames %6d: %d of %d samples off by > 1 LSB (%.1e)"
          % (r, nv, g, sm, e, sc, sr, start, n, int(bad.sum()), bad.size, bad.mean()), flush=True)
    assert bad.mean() < 2e-5, "parity failure"
print("worst share of samples off by more than 1 LSB: %.2e (tests allow 2e-5)" % worst)
