"""
Generates tests/golden/audioop_golden.npz from the REAL CPython stdlib audioop
(Modules/audioop.c; run with CPython 3.12.x — the module is removed in 3.13).
    python tests/golden/make_audioop_golden.py
The vectors pin the integer PCM arithmetic that synthplayer's Sample delegates to; they travel
to the GPU box, where /root/reference and (on newer images) audioop may not exist.
"""
import os
import sys
import warnings

import numpy as np

warnings.simplefilter("ignore", DeprecationWarning)
import audioop  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    rng = np.random.default_rng(20261106)
    a = rng.integers(-32768, 32768, 4099).astype("<i2")
    b = rng.integers(-32768, 32768, 4099).astype("<i2")
    a[:6] = [32767, -32768, -32767, 0, 1, -1]
    b[:6] = [32767, -32768, 32767, 0, -1, -1]
    out = {"a": a, "b": b, "python": np.array(sys.version)}
    A, B = a.tobytes(), b.tobytes()
    out["add"] = np.frombuffer(audioop.add(A, B, 2), "<i2")
    factors = [0.5, 1.5, -1.0, 0.333, 2.75, 0.0, -0.9999, 1e-3]
    out["mul_factors"] = np.array(factors)
    out["mul"] = np.stack([np.frombuffer(audioop.mul(A, 2, f), "<i2") for f in factors])
    biases = [1234, -30000, 40000]
    out["bias_values"] = np.array(biases)
    out["bias"] = np.stack([np.frombuffer(audioop.bias(A, 2, v), "<i2") for v in biases])
    out["reverse"] = np.frombuffer(audioop.reverse(A, 2), "<i2")
    out["max"] = np.array(audioop.max(A, 2))
    st = a[:4098].tobytes()
    out["tomono"] = np.frombuffer(audioop.tomono(st, 2, 0.3, 0.9), "<i2")
    out["tostereo"] = np.frombuffer(audioop.tostereo(A, 2, 0.7, 1.4), "<i2")
    out["lin2lin_2_4"] = np.frombuffer(audioop.lin2lin(A, 2, 4), "<i4")
    out["lin2lin_4_2"] = np.frombuffer(audioop.lin2lin(out["lin2lin_2_4"].tobytes(), 4, 2), "<i2")
    rates = [(44100, 48000), (48000, 44100), (22050, 48000), (48000, 8000), (44100, 44101), (8000, 96000)]
    out["ratecv_rates"] = np.array(rates)
    for i, (ri, ro) in enumerate(rates):
        out["ratecv_mono_%d" % i] = np.frombuffer(audioop.ratecv(A, 2, 1, ri, ro, None)[0], "<i2")
        out["ratecv_stereo_%d" % i] = np.frombuffer(audioop.ratecv(st, 2, 2, ri, ro, None)[0], "<i2")
    np.savez_compressed(os.path.join(HERE, "audioop_golden.npz"), **out)
    print("wrote audioop_golden.npz with", len(out), "arrays from Python", sys.version.split()[0])


if __name__ == "__main__":
    main()
