"""
The C restatement (oracle/oracle.c) is bit-identical to the pure-Python one
(oracle/synth_oracle.py): same IEEE double operations in the same order, same libm sin.
CPU only, small sizes (the Python loops are slow).
"""
import itertools

import numpy as np
import pytest

from oracle import coracle as co
from oracle import synth_oracle as so

SR = 48000
N = 6000
CLS = [so.Sine, so.Triangle, so.Square, so.Sawtooth, so.Pulse]


@pytest.mark.parametrize("wave", range(5))
def test_plain_and_fm_oscillators(wave):
    o = CLS[wave](441.0, 0.8, 0.25, 0.01, samplerate=SR)
    np.testing.assert_array_equal(np.array(so.render_floats(o, N)), co.osc(wave, N, SR, 441.0, 0.8, 0.25, 0.01))
    o = CLS[wave](441.0, 0.8, 0.25, 0.01, fm_lfo=so.Triangle(5.0, 0.05, samplerate=SR), samplerate=SR)
    fm = co.osc(1, N, SR, 5.0, 0.05)
    np.testing.assert_array_equal(np.array(so.render_floats(o, N)), co.osc(wave, N, SR, 441.0, 0.8, 0.25, 0.01, fm=fm))


def test_pulse_pwm_and_negative_phase():
    pwm = co.osc(0, N, SR, 2.0, 0.6, 0.0, 0.5)
    o = so.Pulse(300.0, 0.7, -0.3, 0.0, pulsewidth=0.2, pwm_lfo=so.Sine(2.0, 0.6, bias=0.5, samplerate=SR), samplerate=SR)
    np.testing.assert_array_equal(np.array(so.render_floats(o, N)), co.osc(4, N, SR, 300.0, 0.7, -0.3, 0.0, pwm=pwm, pulsewidth=0.2))
    o = so.Square(300.0, 0.7, -0.3, samplerate=SR)
    np.testing.assert_array_equal(np.array(so.render_floats(o, N)), co.osc(2, N, SR, 300.0, 0.7, -0.3))


def test_noise_injected():
    u = np.random.default_rng(2).random(N)
    o = so.WhiteNoise(12000.0, 0.9, 0.05, samplerate=SR, noise_source=u)
    np.testing.assert_array_equal(np.array(so.render_floats(o, N)), co.noise(N, SR, 12000.0, 0.9, 0.05, u))


@pytest.mark.parametrize("adsr", [(0.01, 0.02, 0.03, 0.6, 0.04), (0.0, 0.02, 0.0, 0.5, 0.03), (0.02, 0.0, 0.01, 1.0, 0.0)])
def test_envelope(adsr):
    src = co.osc(3, N, SR, 440.0)
    for stop in (False, True):
        o = so.EnvelopeFilter(so.Sawtooth(440.0, samplerate=SR), *adsr, stop_at_end=stop)
        np.testing.assert_array_equal(np.array(so.render_floats(o, N)), co.envelope(src, SR, *adsr, stop_at_end=stop))


def test_echo_and_pcm():
    env = (0.01, 0.02, 0.03, 0.6, 0.04)
    o = so.EchoFilter(so.EnvelopeFilter(so.Sawtooth(440.0, 0.5, samplerate=SR), *env), 0.02, 4, 0.013, 0.6)
    want = np.array(so.render_floats(o, N))
    got = co.echo(co.envelope(co.osc(3, N, SR, 440.0, 0.5), SR, *env), SR, 0.02, 4, 0.013, 0.6)
    np.testing.assert_array_equal(got, want)
    pcm, bad = co.to_pcm16(got)
    assert bad == 0 and pcm.tolist() == so.to_pcm(want)
    with pytest.raises(OverflowError):
        so.to_pcm([1.5])
    assert co.to_pcm16(np.array([1.5, -1.5]))[1] == 2


def test_voice_bank_equals_mixing_filter():
    V = 5
    P = np.zeros((V, co.VOICE_STRIDE))
    left, right = [], []
    for v in range(V):
        cw, mw = v % 5, [0, 1, -1, 3, 0][v]
        P[v] = [cw, 200 + 50 * v, 0.15, 0.1 * v, 0.0, 0.3, mw, 3.0 + v, 0.05, 0.2, 0.0,
                0.01, 0.02, 0.03, 0.5, 0.04, 0.3 + 0.1 * v, 0.9 - 0.1 * v]

        def mk():
            fm = CLS[mw](3.0 + v, 0.05, 0.2, 0.0, samplerate=SR) if mw >= 0 else None
            kw = dict(pulsewidth=0.3) if cw == 4 else {}
            return so.EnvelopeFilter(CLS[cw](200 + 50 * v, 0.15, 0.1 * v, 0.0, fm_lfo=fm, samplerate=SR, **kw),
                                     0.01, 0.02, 0.03, 0.5, 0.04)
        left.append(so.AmpModulationFilter(mk(), itertools.repeat(0.3 + 0.1 * v)))
        right.append(so.AmpModulationFilter(mk(), itertools.repeat(0.9 - 0.1 * v)))
    cl, cr = co.voice_bank(P, N, SR)
    np.testing.assert_array_equal(np.array(so.render_floats(so.MixingFilter(*left), N)), cl)
    np.testing.assert_array_equal(np.array(so.render_floats(so.MixingFilter(*right), N)), cr)


def test_blocks_and_generator_shapes():
    o = so.Sine(440, samplerate=SR)
    b = list(itertools.islice(o.blocks(), 3))
    assert [len(x) for x in b] == [512, 512, 512]
    assert list(itertools.islice(o.generator(), 5)) == b[0][:5]
    e = so.EnvelopeFilter(so.Sine(440, samplerate=SR), 0.001, 0.001, 0.001, 0.5, 0.001, stop_at_end=True)
    total = sum(len(x) for x in e.blocks())
    assert 190 <= total <= 194
