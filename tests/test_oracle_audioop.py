"""
Pins the integer-PCM part of the oracle: oracle/audioop_port.py and the C restatement
(oracle/oracle.c) against (1) the committed golden vectors generated from the real CPython
audioop and (2) the real module itself when importable.  CPU only.
"""
import os
import warnings

import numpy as np
import pytest

from oracle import audioop_port as port
from oracle import coracle as co

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "audioop_golden.npz"))

try:
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", DeprecationWarning)
        import audioop as real
except ImportError:          # Python >= 3.13
    real = None


def i2(b):
    return np.frombuffer(b, "<i2")


def test_port_matches_golden():
    A, B = G["a"].tobytes(), G["b"].tobytes()
    st = G["a"][:4098].tobytes()
    np.testing.assert_array_equal(i2(port.add(A, B, 2)), G["add"])
    for f, want in zip(G["mul_factors"], G["mul"]):
        np.testing.assert_array_equal(i2(port.mul(A, 2, float(f))), want)
    for v, want in zip(G["bias_values"], G["bias"]):
        np.testing.assert_array_equal(i2(port.bias(A, 2, int(v))), want)
    np.testing.assert_array_equal(i2(port.reverse(A, 2)), G["reverse"])
    assert port.max(A, 2) == int(G["max"])
    np.testing.assert_array_equal(i2(port.tomono(st, 2, 0.3, 0.9)), G["tomono"])
    np.testing.assert_array_equal(i2(port.tostereo(A, 2, 0.7, 1.4)), G["tostereo"])
    np.testing.assert_array_equal(np.frombuffer(port.lin2lin(A, 2, 4), "<i4"), G["lin2lin_2_4"])
    np.testing.assert_array_equal(i2(port.lin2lin(G["lin2lin_2_4"].tobytes(), 4, 2)), G["lin2lin_4_2"])


def test_ratecv_port_closed_form_and_c_match_golden():
    A = G["a"].tobytes()
    st = G["a"][:4098].tobytes()
    for i, (ri, ro) in enumerate(G["ratecv_rates"]):
        ri, ro = int(ri), int(ro)
        want_m, want_s = G["ratecv_mono_%d" % i], G["ratecv_stereo_%d" % i]
        np.testing.assert_array_equal(i2(port.ratecv(A, 2, 1, ri, ro, None)[0]), want_m)
        np.testing.assert_array_equal(i2(port.ratecv_closed_form(A, 2, 1, ri, ro)), want_m)
        np.testing.assert_array_equal(i2(port.ratecv_closed_form(st, 2, 2, ri, ro)), want_s)
        np.testing.assert_array_equal(co.pcm16_ratecv(G["a"], 1, ri, ro), want_m)
        np.testing.assert_array_equal(co.pcm16_ratecv(G["a"][:4098], 2, ri, ro), want_s)


def test_c_port_matches_golden():
    np.testing.assert_array_equal(co.pcm16_add(G["a"], G["b"]), G["add"])
    for f, want in zip(G["mul_factors"], G["mul"]):
        np.testing.assert_array_equal(co.pcm16_mul(G["a"], float(f)), want)


@pytest.mark.skipif(real is None, reason="stdlib audioop removed in this Python")
def test_port_matches_live_audioop_on_random_buffers():
    rng = np.random.default_rng(1)
    for trial in range(20):
        n = int(rng.integers(1, 3000)) * 2
        a = rng.integers(-32768, 32768, n).astype("<i2").tobytes()
        b = rng.integers(-32768, 32768, n).astype("<i2").tobytes()
        f = float(rng.uniform(-3, 3))
        assert port.add(a, b, 2) == real.add(a, b, 2)
        assert port.mul(a, 2, f) == real.mul(a, 2, f)
        assert port.bias(a, 2, int(rng.integers(-70000, 70000))) is not None
        assert port.tomono(a, 2, f, 1 - f) == real.tomono(a, 2, f, 1 - f)
        assert port.tostereo(a, 2, f, 1 - f) == real.tostereo(a, 2, f, 1 - f)
        assert port.max(a, 2) == real.max(a, 2)
        assert port.minmax(a, 2) == real.minmax(a, 2)
        assert port.rms(a, 2) == real.rms(a, 2)
        assert port.reverse(a, 2) == real.reverse(a, 2)
        assert port.lin2lin(a, 2, 4) == real.lin2lin(a, 2, 4)
        ri, ro = int(rng.integers(4000, 100000)), int(rng.integers(4000, 100000))
        for ch in (1, 2):
            want = real.ratecv(a, 2, ch, ri, ro, None)[0]
            assert port.ratecv_closed_form(a, 2, ch, ri, ro) == want
        bias = int(rng.integers(-70000, 70000))
        assert port.bias(a, 2, bias) == real.bias(a, 2, bias)


def test_ratecv_edge_cases():
    assert port.ratecv_closed_form(b"", 2, 1, 44100, 48000) == b""
    one = np.array([1234], "<i2").tobytes()
    assert i2(port.ratecv_closed_form(one, 2, 1, 44100, 48000)).tolist() == i2(port.ratecv(one, 2, 1, 44100, 48000, None)[0]).tolist()


def test_sample_oracle_fades_match_c():
    from oracle.sample_oracle import Sample
    rng = np.random.default_rng(3)
    a = rng.integers(-32768, 32768, 5000).astype("<i2")
    s = Sample.from_raw_frames(a.tobytes(), 2, 10000, 1)
    np.testing.assert_array_equal(i2(s.copy().fadeout(0.5, 0.2)._frames), co.pcm16_fadeout(a, 0.2))
    np.testing.assert_array_equal(i2(s.copy().fadein(0.5, 0.1)._frames), co.pcm16_fadein(a, 0.1))
