"""
GPU parity of the fused voice-bank kernel and of BASELINE.json's configurations against the
C oracle (serial fp64).  Tolerance +-1 LSB of int16 PCM (north_star).
"""
import ctypes

import numpy as np
import pytest

from helpers import assert_within_lsb, dev_pcm16, pcm16_of

pytestmark = pytest.mark.gpu


def make_bank(nvoices, sr, seed, duration, waves=None, fm=True):
    """Random bank; returns (VoiceBank, oracle parameter matrix)."""
    import synthesizer_b200 as sb
    rng = np.random.default_rng(seed)
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    bank = sb.VoiceBank(sr)
    P = np.zeros((nvoices, 18))
    for v in range(nvoices):
        cw = int(rng.integers(0, 5)) if waves is None else waves[v % len(waves)]
        mw = int(rng.integers(0, 5)) if fm and rng.random() < 0.8 else -1
        cf = float(rng.uniform(55, 1760))
        mf = float(rng.uniform(0.5, 12))
        depth = float(rng.uniform(0.005, 0.06))
        amp = float(rng.uniform(0.2, 1.0)) / nvoices
        a, d, r = (float(x) for x in rng.uniform(0.01, 0.2 * duration, 3))
        s = max(0.0, duration * float(rng.uniform(0.3, 0.9)) - a - d - r)
        sl = float(rng.uniform(0.2, 1.0))
        pan = float(rng.uniform(-1, 1))
        gl, gr = (1 - pan) / 2, (1 + pan) / 2
        cph, mph = float(rng.random()), float(rng.random())
        cpw, mpw = float(rng.uniform(0.1, 0.9)), float(rng.uniform(0.1, 0.9))
        lfo = None
        if mw >= 0:
            kw = dict(pulsewidth=mpw) if mw == 4 else {}
            lfo = classes[mw](mf, depth, mph, 0.0, samplerate=sr, **kw)
        kw = dict(pulsewidth=cpw) if cw == 4 else {}
        car = classes[cw](cf, amp, cph, 0.0, fm_lfo=lfo, samplerate=sr, **kw)
        bank.add(sb.EnvelopeFilter(car, a, d, s, sl, r), gl, gr)
        # the C struct stores gains and amplitudes as float32: give the oracle the same values
        f32 = lambda x: float(np.float32(x))
        P[v] = [cw, cf, f32(amp), cph, 0.0, cpw, mw, mf, depth, mph, 0.0, a, d, s, sl, r, f32(gl), f32(gr)]
        if mw == 4:
            pass
    return bank, P


def force_some_generic(bank, P):
    """Gives every other voice with a non-sine modulator a negative modulator start phase: the
    fast path's periodic rules need it non-negative, so those voices take the generic kernel
    (negative phases are legal upstream: Python's % and int() define the waveform there)."""
    k = 0
    for i, v in enumerate(bank._voices):
        if v.m_wave > 0:
            k += 1
            if k % 2:
                v.m_phase = P[i, 9] = -0.3 - 0.01 * i
    bank._dev = None


def oracle_bank_pcm(P, nframes, sr, mpw=None):
    from oracle import coracle as co
    left, right = co.voice_bank(P, nframes, sr)
    out = np.empty(2 * nframes, dtype=np.int64)
    out[0::2] = pcm16_of(left)
    out[1::2] = pcm16_of(right)
    return out


@pytest.mark.parametrize("nvoices,seconds", [(1, 1.0), (7, 0.5), (64, 0.25), (130, 0.3)])
def test_bank_matches_oracle(cuda, nvoices, seconds):
    sr = 48000
    n = int(sr * seconds) + 37
    # modulator pulse widths other than the oracle's fixed 0.1 are not expressible there
    bank, P = make_bank(nvoices, sr, 40 + nvoices, seconds, fm=True)
    for v in bank._voices:
        if v.m_wave == 4:
            v.m_pw = 0.1
    got = dev_pcm16(bank.render(n))
    want = oracle_bank_pcm(P, n, sr)
    bad = np.abs(got.astype(np.int64) - want) > 1
    # FM'd discontinuous carriers may flip an edge sample (fp64 noise of the serial recurrence);
    # each voice is 1/nvoices of full scale, so such a flip is visible but bounded
    assert bad.mean() < 2e-5, "%d of %d samples off by more than 1 LSB" % (bad.sum(), bad.size)


def test_bank_sine_voices_exact_tolerance(cuda):
    """Only continuous waveforms (sine, triangle): every sample within +-1 LSB."""
    sr, seconds = 48000, 2.0
    n = int(sr * seconds)
    bank, P = make_bank(32, sr, 77, seconds, waves=[0, 1], fm=True)
    for i, v in enumerate(bank._voices):
        if v.m_wave in (2, 3, 4):       # keep modulators continuous too
            v.m_wave = 0
            P[i, 6] = 0
    got = dev_pcm16(bank.render(n))
    assert_within_lsb(got, oracle_bank_pcm(P, n, sr), 1, "sine/triangle bank")


def test_fast_path_ten_minutes_in(cuda):
    """The closed-form FM sums and the exact time base must still agree with upstream's serial
    recurrence 10 minutes into a render (the time base has crossed ~25 binades by then and
    both the carrier's and the modulator's accumulated t have drifted)."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    sr = 44100
    start, n = 600 * sr - 6000, 6000
    specs = [  # (carrier wave, f0, mod freq, depth, mod bias)
        (0, 2793.8, 6.1, 0.04, 0.0), (0, 440.0, 0.13, 0.3, 0.0), (1, 987.7, 110.0, 0.02, 0.01), (0, 61.7, 880.0, 1.5, 0.0)]
    bank = sb.VoiceBank(sr)
    P = np.zeros((len(specs), 18))
    f32 = lambda x: float(np.float32(x))
    for i, (cw, f0, fm, d, mb) in enumerate(specs):
        lfo = sb.Sine(fm, f32(d), 0.3, f32(mb), samplerate=sr)
        car = (sb.Sine if cw == 0 else sb.Triangle)(f0, f32(0.22), 0.1 * i, 0.0, fm_lfo=lfo, samplerate=sr)
        bank.add(car, f32(0.9 - 0.2 * i), f32(0.2 + 0.2 * i))
        P[i] = [cw, f0, f32(0.22), 0.1 * i, 0, 0.1, 0, fm, f32(d), 0.3, f32(mb), -1, 0, 0, 0, 0, f32(0.9 - 0.2 * i), f32(0.2 + 0.2 * i)]
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[i]), float(sr)) == 1
    got = dev_pcm16(bank.render(n, start=start))
    from oracle import coracle as co
    left, right = co.voice_bank(P, start + n, sr)
    want = np.empty(2 * n, dtype=np.int64)
    want[0::2] = pcm16_of(left[start:])
    want[1::2] = pcm16_of(right[start:])
    assert_within_lsb(got, want, 1, "fast path at 10 minutes")


def test_edged_carriers_ten_minutes_in(cuda):
    """Square / Sawtooth / Pulse carriers without FM, 10 minutes into a render, one voice at a
    time at full scale: every edge must still fall on the sample upstream puts it on — also at
    frequencies that divide the sample rate, where all edges sit on sample boundaries and the
    side is decided by the rounding of the accumulated t, and at a high frequency, where the
    16-sample integer-phase run drifts furthest from the time base's own step."""
    import synthesizer_b200 as sb
    from oracle import coracle as co
    sr = 44100
    start, n = 600 * sr - 20000, 20000
    classes = {2: sb.Square, 3: sb.Sawtooth, 4: sb.Pulse}
    for cw, f0, ph in [(2, 2205.0, 0.0), (2, 441.0, 0.25), (2, 9973.3, 0.1), (3, 1102.5, 0.5), (3, 15011.7, 0.3),
                       (4, 4410.0, 0.0), (4, 733.1, 0.9)]:
        kw = dict(pulsewidth=0.3) if cw == 4 else {}
        car = classes[cw](f0, 0.99, ph, 0.0, samplerate=sr, **kw)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert bank.count_generic() == 0 and bank.count_edged() == 1
        got = bank.render(n, start=start).cpu().numpy()[0::2].astype(np.float64)
        want = co.osc(cw, start + n, sr, f0, float(np.float32(0.99)), ph, 0.0, pulsewidth=0.3)[start:]
        err = np.abs(got - want)
        assert err.max() < 1e-6, "wave %d f0 %g: %d samples differ, first at %d" % (
            cw, f0, int((err >= 1e-6).sum()), int(np.argmax(err >= 1e-6)))


@pytest.mark.parametrize("cw", [0, 1, 3])
def test_fm_accuracy_sweep_full_scale(cuda, cw):
    """One FULL-SCALE voice per (carrier f0, modulator f, FM depth) — from slow deep vibrato to
    audio-rate FM with index 2 on a 10 kHz carrier — against the oracle.  The error budget of
    +-1 LSB has to hold at full amplitude, on whichever kernel class the voice lands.
    Frequencies are chosen incommensurate with the sample rate: with 3520 Hz x 2000 Hz at 48 kHz
    the sawtooth's exact phase comes within 1e-13..1e-9 cycles of a wrap every 1200 samples, and
    on which side such a sample falls is decided by the rounding noise of upstream's own fp64
    recurrence — no other evaluation order can reproduce it."""
    import itertools
    import synthesizer_b200 as sb
    from oracle import coracle as co
    sr, n = 48000, 48000
    classes = {0: sb.Sine, 1: sb.Triangle, 3: sb.Sawtooth}
    seen = set()
    # (depth 5 on the 10 kHz carrier is a gain of 6.5 rad: generic for every carrier; depth 2 there, 2.6 rad,
    # is the strong-FM variant of the fast path for Sine / Triangle carriers and generic for Sawtooth)
    for f0, fm, d in itertools.product((55.1, 441.3, 3517.9, 9973.0), (0.1, 5.3, 109.7, 1999.3), (0.01, 0.5, 2.0, 5.0)):
        lfo = sb.Sine(fm, d, 0.3, samplerate=sr)
        car = classes[cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        seen.add(bank.count_generic())
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, fm, d, 0.3, 0.0)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        flips = err > 100          # a sawtooth wrap one sample early/late (fp64 noise of upstream's recurrence)
        assert flips.sum() <= (3 if cw == 3 else 0), (f0, fm, d, int(flips.sum()))
        assert err[~flips].max() < 0.9, "f0=%g fm=%g depth=%g: %.3f LSB" % (f0, fm, d, err[~flips].max())
    assert seen == {0, 1}, "the sweep is meant to exercise both the fast and the generic kernels"


@pytest.mark.parametrize("cw", [0, 1])
def test_strong_fm_stays_on_the_fast_path(cuda, cw):
    """FM gains between 1.5 and 6 rad per unit of modulator sum (a bright FM patch in a high octave:
    f0 * depth of 11 - 46 kHz): Sine / Triangle carriers with a Sine modulator take the fast path's
    strong-FM variant (phase reduced after every step).  Full-scale voices against the oracle, the
    error bound the plain run is held to (half an LSB), plus a bank that mixes them with plain voices."""
    import itertools
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, n = 48000, 96000
    classes = {0: sb.Sine, 1: sb.Triangle}
    for gain, f0, fm in itertools.product((1.6, 3.0, 4.5, 5.9), (1979.3, 7013.9), (2.7, 219.3, 3001.7)):
        d = gain * sr / (2 * np.pi * f0)
        lfo = sb.Sine(fm, d, 0.3, samplerate=sr)
        car = classes[cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 1, (gain, f0, fm)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, fm, d, 0.3, 0.0)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        assert err.max() < 0.5, "cw=%d gain=%g f0=%g fm=%g: %.3f LSB" % (cw, gain, f0, fm, err.max())
    # ... and with the piecewise-linear modulators (class 2: their values are formed exactly per run)
    mods = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    for mw, gain, f0, fm in itertools.product((1, 2, 3, 4), (2.5, 5.9), (3979.3,), (3.7, 431.9)):
        d = gain * sr / (2 * np.pi * f0)
        mkw = dict(pulsewidth=0.37) if mw == 4 else {}
        lfo = mods[mw](fm, d, 0.3, 0.0, samplerate=sr, **mkw)
        car = classes[cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 2, (mw, gain, f0, fm)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(mw, n, sr, fm, d, 0.3, 0.0, pulsewidth=0.37)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        assert err.max() < 0.5, "cw=%d mw=%d gain=%g fm=%g: %.3f LSB" % (cw, mw, gain, fm, err.max())
    # just above the limit: generic
    d = 6.1 * sr / (2 * np.pi * 7013.9)
    car = classes[cw](7013.9, 0.5, fm_lfo=sb.Sine(5.0, d, samplerate=sr), samplerate=sr)
    assert sb.VoiceBank(sr).add(car, 1.0, 1.0).count_generic() == 1
    # a bank of 100 voices, every third one strong: pairs of plain voices, strong ones taken alone
    bank, P = make_bank(100, sr, 123 + cw, 0.3, waves=[cw], fm=True)
    for i, v in enumerate(bank._voices):
        if v.m_wave > 0:
            v.m_wave = P[i, 6] = 0
        if i % 3 == 0 and v.m_wave == 0:
            v.c_freq = P[i, 1] = 3000.0 + 17.0 * i
            v.m_amp = P[i, 8] = float(np.float32(4.0 * sr / (2 * np.pi * v.c_freq)))
    bank._dev = None
    assert bank.count_generic() == 0 and sum(1 for i, v in enumerate(bank._voices) if i % 3 == 0 and v.m_wave == 0) > 15
    nn = 14000
    got = dev_pcm16(bank.render(nn))
    assert_within_lsb(got, oracle_bank_pcm(P, nn, sr), 1, "bank with strong-FM voices")


@pytest.mark.parametrize("pw", [0.0, 1.0, 1e-12, 1.0 - 1e-12, 0.5, 0.999999])
def test_pulse_carrier_extreme_widths(cuda, pw):
    """Pulse widths at and beyond the ends of (0, 1) — upstream clamps them to one ulp inside —
    on the integer-phase path, where a width of almost 1 rounds to 2^32."""
    import synthesizer_b200 as sb
    from oracle import coracle as co
    sr, n = 44100, 30000
    for f0, fmlfo in ((441.0, None), (997.3, sb.Sine(5.0, 0.03, samplerate=sr))):
        car = sb.Pulse(f0, 0.99, 0.2, 0.0, pulsewidth=pw, fm_lfo=fmlfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert bank.count_generic() == 0
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, 5.0, 0.03, 0.0, 0.0) if fmlfo is not None else None
        want = co.osc(4, n, sr, f0, float(np.float32(0.99)), 0.2, 0.0, fm=m, pulsewidth=pw)
        flips = np.abs(got - want) > 1e-6
        assert flips.sum() <= (2 if fmlfo is not None else 0), (pw, f0, int(flips.sum()))


def test_sine_modulator_corner_cases(cuda):
    """Closed-form class at its edges: modulator steps close to the bounds the classifier
    admits (1e-7 < |w| < 3 rad per sample), a negative modulator frequency, a modulator at a
    start phase above one cycle; and just outside them, where the generic kernel takes over."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, n = 48000, 60000
    for fm, d, mph, want_cls in [(0.002, 0.4, 0.25, 1), (22000.0, 0.3, 0.0, 1), (-6.3, 0.2, 0.7, 1), (5.1, 0.3, 3.75, 1),
                                 (0.0005, 0.4, 0.25, 0), (23500.0, 0.3, 0.0, 0)]:
        lfo = sb.Sine(fm, d, mph, 0.0, samplerate=sr)
        car = sb.Sine(1318.5, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == want_cls, (fm, d)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, fm, d, mph, 0.0)
        want = co.osc(0, n, sr, 1318.5, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        assert err.max() < 0.9, "modulator f=%g depth=%g phase=%g: %.3f LSB" % (fm, d, mph, err.max())


def test_piecewise_modulator_corner_cases(cuda):
    """Tile-scan class at its edges: a modulator that never moves (0 Hz), one at Nyquist, zero
    depth, and Pulse modulators whose width is clamped at either end."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, n = 48000, 40000
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    cases = [(1, 0.0, 0.3, 0.1), (2, 24000.0, 0.2, 0.1), (3, 7.7, 0.0, 0.1), (4, 3.3, 0.25, 0.0), (4, 3.3, 0.25, 1.0),
             (2, 0.0, 0.5, 0.1), (4, 12000.0, 0.1, 0.5)]
    for mw, fm, d, mpw in cases:
        kw = dict(pulsewidth=mpw) if mw == 4 else {}
        lfo = classes[mw](fm, d, 0.6, 0.0, samplerate=sr, **kw)
        car = sb.Sine(523.3, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 2, (mw, fm, d, mpw)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(mw, n, sr, fm, d, 0.6, 0.0, pulsewidth=mpw)
        want = co.osc(0, n, sr, 523.3, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        assert err.max() < 0.9, "modulator %d f=%g depth=%g pw=%g: %.3f LSB" % (mw, fm, d, mpw, err.max())


@pytest.mark.parametrize("cw", [2, 3, 4])
def test_edged_carriers_with_modulator_bias(cuda, cw):
    """A modulator bias scales the carrier frequency (f0*(1+bias)): also above Nyquist, where the
    integer phase wraps more than once per sample, and below zero, where the phase runs
    backwards and the periodic rules for the edges no longer hold (exact path)."""
    import synthesizer_b200 as sb
    from oracle import coracle as co
    sr, n = 48000, 48000
    classes = {2: sb.Square, 3: sb.Sawtooth, 4: sb.Pulse}
    for f0, bias in ((441.3, 0.3), (2217.7, 20.5), (661.1, -1.5), (97.3, -0.999)):
        lfo = sb.Sine(3.1, 0.02, 0.1, bias, samplerate=sr)
        kw = dict(pulsewidth=0.3) if cw == 4 else {}
        car = classes[cw](f0, 0.99, 0.2, 0.0, fm_lfo=lfo, samplerate=sr, **kw)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, 3.1, 0.02, 0.1, bias)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.2, 0.0, fm=m, pulsewidth=0.3)
        err = np.abs(got - want)
        flips = err > 1e-3
        # (a sawtooth far above Nyquist steps by most of its range per sample: its fp32 ramp
        # value carries the phase error of the run, 2e-5 cycles at worst)
        assert flips.sum() <= 3, (cw, f0, bias, int(flips.sum()), bank.count_generic())
        assert err[~flips].max() < 1e-4, (cw, f0, bias, float(err[~flips].max()))


@pytest.mark.parametrize("cw", [0, 1, 3])
def test_deep_fm_on_low_carriers(cuda, cw):
    """Depths well above 1 (the instantaneous frequency goes negative, the phase runs backwards
    for part of every modulator cycle) stay on the fast path as long as the FM gain
    2*pi*f0*depth/sr is at most 1.5 rad: full-scale voices against the oracle."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, n = 48000, 96000
    classes = {0: sb.Sine, 1: sb.Triangle, 3: sb.Sawtooth}
    for f0, fm, d in [(55.1, 5.3, 20.0), (110.3, 109.7, 12.0), (441.3, 0.13, 6.0), (27.5, 880.3, 60.0)]:
        lfo = sb.Sine(fm, d, 0.3, samplerate=sr)
        car = classes[cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 1, (f0, fm, d)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, fm, d, 0.3, 0.0)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        flips = err > 100
        assert flips.sum() <= (3 if cw == 3 else 0), (f0, fm, d, int(flips.sum()))
        assert err[~flips].max() < 0.9, "cw=%d f0=%g fm=%g depth=%g: %.3f LSB" % (cw, f0, fm, d, err[~flips].max())


@pytest.mark.parametrize("cw", [2, 4])
def test_deep_fm_pulls_an_edged_carrier_below_zero(cuda, cw):
    """Square / Pulse carriers whose FM index is above one and whose modulator starts on its
    negative half: the phase dips below zero in the first frames, where Square's int() is not the
    periodic rule - those runs are evaluated exactly, and only those (the rest of the render is on
    the integer-phase path)."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, n = 48000, 48000
    classes = {2: sb.Square, 4: sb.Pulse}
    for f0, fm, d, mph in [(55.1, 5.3, 6.0, 0.6), (110.3, 2.1, 3.0, 0.75), (27.5, 40.3, 20.0, 0.55), (220.7, 1.3, 1.01, 0.5)]:
        lfo = sb.Sine(fm, d, mph, samplerate=sr)
        kw = dict(pulsewidth=0.3) if cw == 4 else {}
        car = classes[cw](f0, 0.99, 0.0, 0.0, fm_lfo=lfo, samplerate=sr, **kw)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 1, (f0, fm, d)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(0, n, sr, fm, d, mph, 0.0)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.0, 0.0, fm=m, pulsewidth=0.3)
        flips = np.abs(got - want) > 1e-3
        assert flips.sum() <= 3, (cw, f0, fm, d, int(flips.sum()), np.flatnonzero(flips)[:8])


@pytest.mark.parametrize("mw", [1, 2, 3, 4])
def test_piecewise_modulators_across_a_late_seam_of_the_time_base(cuda, mw):
    """Class-2 voices rendered across t = 256 s, where the serial time base's step changes: past
    128 s a run that crosses such a seam takes the exact per-sample path (before it, the straight
    line holds across seams).  Modulators slow and fast enough to put one and several edges in a run."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr = 48000
    start, n = 256 * sr - 3000, 6000
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    for cw, f0, fm, d in [(0, 441.3, 6.1, 0.2), (3, 1103.7, 331.3, 0.03), (2, 97.3, 2903.1, 0.01), (1, 661.1, 0.37, 0.4)]:
        mkw = dict(pulsewidth=0.37) if mw == 4 else {}
        lfo = classes[mw](fm, d, 0.3, 0.0, samplerate=sr, **mkw)
        car = classes[cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 2, (cw, f0, fm, d)
        got = bank.render(n, start=start).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(mw, start + n, sr, fm, d, 0.3, 0.0, pulsewidth=0.37)
        want = co.osc(cw, start + n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)[start:]
        err = np.abs(got - want) * 32767.0
        flips = err > 100
        assert flips.sum() <= (3 if cw >= 2 else 0), (cw, f0, fm, d, int(flips.sum()))
        assert err[~flips].max() < 0.9, "mw=%d cw=%d f0=%g fm=%g: %.3f LSB" % (mw, cw, f0, fm, err[~flips].max())


@pytest.mark.parametrize("mw", [1, 2, 3, 4])
def test_piecewise_modulators_full_scale(cuda, mw):
    """Triangle / Square / Sawtooth / Pulse modulators on the fast path (class 2: exact fp64
    modulator values per run, tile pre-pass + warp scan for the FM sum): one FULL-SCALE voice per
    (carrier, f0, modulator f, depth) against the oracle, 3 s so that the sum runs over many
    tiles, and a second render started mid-way that must continue the same sum."""
    import itertools
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, n, cut = 48000, 144000, 70001
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    for cw, f0, fm, d in itertools.product((0, 3), (110.3, 2217.7), (0.37, 6.1, 331.3), (0.02, 0.4)):
        mkw = dict(pulsewidth=0.37) if mw == 4 else {}
        lfo = classes[mw](fm, d, 0.3, 0.05, samplerate=sr, **mkw)
        car = classes[cw](f0, 0.99, 0.1, 0.0, fm_lfo=lfo, samplerate=sr)
        bank = sb.VoiceBank(sr).add(car, 1.0, 1.0)
        assert C.lib.sb_voice_is_fast(ctypes.byref(bank._voices[0]), float(sr)) == 2, (cw, f0, fm, d)
        got = bank.render(n).cpu().numpy()[0::2].astype(np.float64)
        m = co.osc(mw, n, sr, fm, d, 0.3, 0.05, pulsewidth=0.37)
        want = co.osc(cw, n, sr, f0, float(np.float32(0.99)), 0.1, 0.0, fm=m)
        err = np.abs(got - want) * 32767.0
        flips = err > 100          # a sawtooth carrier's wrap one sample early/late
        assert flips.sum() <= (3 if cw == 3 else 0), (cw, f0, fm, d, int(flips.sum()))
        assert err[~flips].max() < 0.9, "cw=%d f0=%g fm=%g depth=%g: %.3f LSB" % (cw, f0, fm, d, err[~flips].max())
        tail = bank.render(n - cut, start=cut).cpu().numpy()[0::2].astype(np.float64)
        terr = np.abs(tail - want[cut:]) * 32767.0
        assert (terr > 100).sum() <= (3 if cw == 3 else 0) and terr[terr <= 100].max() < 0.9, (cw, f0, fm, d)


def test_fast_and_generic_voices_share_a_group(cuda):
    """A group holding both classes: the generic kernel adds into what the fast one wrote."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    sr, n = 48000, 20000
    bank, P = make_bank(24, sr, 91, 0.4, fm=True)
    for v in bank._voices:
        if v.m_wave == 4:
            v.m_pw = 0.1
    force_some_generic(bank, P)
    kinds = [C.lib.sb_voice_is_fast(ctypes.byref(v), float(sr)) for v in bank._voices]
    assert set(kinds) == {0, 1, 2}, "the bank is meant to hold generic, closed-form and tile-scan voices"
    assert bank.count_generic() == kinds.count(0) and bank.count_scan() == kinds.count(2)
    got = dev_pcm16(bank.render(n))
    bad = np.abs(got.astype(np.int64) - oracle_bank_pcm(P, n, sr)) > 1
    assert bad.mean() < 2e-5
    # the counts are hints: without them (-1 = unknown: every kernel runs, scratch sized for the
    # worst case) the render is the same, bit for bit
    hinted = bank.render(n).cpu().numpy()
    vd, _ = bank._upload(0, None)
    unknown = bank.render(n, voices_dev=vd).cpu().numpy()
    assert np.array_equal(hinted, unknown)
    # ... and a scan_voices hint that is too small hands the surplus tile-scan voices to the
    # generic kernels (which must then be allowed to run): still within tolerance of the oracle
    short = bank.render(n, voices_dev=vd, generic_voices=len(kinds), edged_voices=len(kinds), scan_voices=1)
    bad = np.abs(dev_pcm16(short).astype(np.int64) - oracle_bank_pcm(P, n, sr)) > 1
    assert bad.mean() < 2e-5
    assert bank.check_last_render() == kinds.count(0) + kinds.count(2) - 1
    # hints that are too small while generic_voices says 0 lose voices: the status call reports it
    bank.render(n, voices_dev=vd, generic_voices=0, edged_voices=len(kinds), scan_voices=len(kinds))
    with pytest.raises(C.SbError, match="out of the mix"):
        bank.check_last_render()
    bank.render(n, voices_dev=vd)
    assert bank.check_last_render() == kinds.count(0)


@pytest.mark.parametrize("waves,sine_mods", [((2, 3, 4), True), ((0, 1, 3), False), ((2, 3, 4), False)])
def test_sorted_bank_without_a_smooth_part(cuda, waves, sine_mods):
    """A bank of several groups whose first kernel part is not part 0: edged carriers only (part 1
    writes the partial mixes), class-2 modulators only (part 2 writes them), or both, each with
    generic voices mixed in - whose groups, at the end of the sorted bank, hold no voice of the
    fast classes and must come out of the first part as zeros for the generic kernel to add into."""
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    sr, n = 48000, 12000
    nv = 200
    bank, P = make_bank(nv, sr, 7 + len(waves) + int(sine_mods), 0.25, waves=list(waves), fm=True)
    for i, v in enumerate(bank._voices):
        if v.m_wave == 4:
            v.m_pw = 0.1
        if sine_mods and v.m_wave > 0:                   # Sine modulators only: no class-2 voice
            v.m_wave = P[i, 6] = 0
        if not sine_mods and v.m_wave <= 0:              # piecewise-linear modulators only: every voice class 2
            v.m_wave = P[i, 6] = 1 + i % 3
            v.m_freq, v.m_amp, v.m_phase = P[i, 7], P[i, 8], P[i, 9] = 3.0 + 0.01 * i, float(np.float32(0.02)), 0.25
    bank._dev = None
    # every fifth voice to the generic kernels (a negative carrier start phase for edged carriers, a
    # negative modulator phase for class-2 modulators)
    for i, v in enumerate(bank._voices):
        if i % 5 == 0:
            if v.m_wave > 0:
                v.m_phase = P[i, 9] = -0.3 - 0.001 * i
            elif v.c_wave >= 2:
                v.c_phase = P[i, 3] = -0.2 - 0.001 * i
    bank._dev = None
    g, sm, e, sc = bank.count_classes()
    assert g > 0 and (e > 0 or sc > 0)
    if waves == (2, 3, 4) and sine_mods:
        assert sm == 0 and sc == 0                       # part 1 is the first part
    if not sine_mods:
        assert sm == 0 and e == 0                        # part 2 is the first part
    got = dev_pcm16(bank.render(n))
    want = oracle_bank_pcm(P, n, sr)
    bad = np.abs(got.astype(np.int64) - want) > 1
    assert bad.mean() < 2e-5, "%d of %d samples off by more than 1 LSB (generic %d smooth %d edged %d scan %d)" % (
        bad.sum(), bad.size, g, sm, e, sc)


@pytest.mark.parametrize("nvoices,hints", [(300, True), (300, False), (70, True)])
def test_render_stays_inside_the_scratch_the_plan_asks_for(cuda, nvoices, hints):
    """sb_bank_render_f32 with scratch, partial-mix and output buffers of exactly the sizes the plan
    states, each followed by a guard area: a mixed bank (all kernel parts, generic voices, the sort,
    the unit queues, the modulator tile table) must leave the guards untouched - with exact hints
    and with none (worst-case sizes)."""
    import torch
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C, _device as D
    sr, n, start = 48000, 30011, 517
    bank, P = make_bank(nvoices, sr, 77 + nvoices, 0.5, fm=True)
    force_some_generic(bank, P)
    vd, nv = bank._upload(0, None)
    g, sm, e, sc = bank.count_classes()
    assert g > 0 and sm > 0 and e > 0 and sc > 0
    plan = C.BankPlan()
    tr, tc = D.timebase_rad(sr), D.timebase_cyc(sr)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    C.call("sb_bank_plan_make", nv, n, start, sms, tr.nsegs, ctypes.byref(plan))
    if hints:
        C.call("sb_bank_plan_hint", ctypes.byref(plan), g, sm, e, sc)
    guard = 1 << 16

    def guarded(nbytes):
        t = torch.full((nbytes + guard,), 0x5A, dtype=torch.uint8, device="cuda")
        return t

    scratch, partial, mix = guarded(plan.carry_bytes), guarded(max(plan.partial_bytes, 16)), guarded(2 * n * 4)
    C.call("sb_bank_render_f32", D.ptr(vd), ctypes.byref(plan), D.ptr(tr.dev()), tr.nsegs, D.ptr(tc.dev()), tc.nsegs,
           start, float(sr), D.ptr(scratch), D.ptr(partial), D.ptr(mix), D.stream())
    torch.cuda.synchronize()
    assert bool((scratch[plan.carry_bytes:] == 0x5A).all()), "scratch overrun"
    assert bool((partial[max(plan.partial_bytes, 16):] == 0x5A).all()), "partial-mix overrun"
    assert bool((mix[2 * n * 4:] == 0x5A).all()), "output overrun"
    # and the mix is the one the product API renders
    want = bank.render(n, start=start)
    got = mix[:2 * n * 4].view(torch.float32)
    assert torch.equal(got, want)


def test_bank_chunked_start_offsets(cuda):
    """Rendering [0,n) in one call equals rendering it as two calls with start offsets
    (non-FM voices; FM voices carry state across calls only through the general path)."""
    sr, n = 44100, 30000
    bank, P = make_bank(16, sr, 5, 0.6, fm=False)
    whole = bank.render(n).cpu().numpy()
    a = bank.render(12345).cpu().numpy()
    b = bank.render(n - 12345, start=12345).cpu().numpy()
    # the split moves every frame to a different lane of the fp32 run arithmetic: equal to
    # fp32 rounding (1e-7 of full scale = 0.003 LSB), not bit for bit
    np.testing.assert_allclose(np.concatenate([a, b]), whole, rtol=0, atol=1e-7)


def test_bank_start_offset_with_fm_voices_of_both_classes(cuda):
    """A render that starts at n0 > 0 must continue every voice's FM sum from frame 0: fast
    voices by closed form, tile-scan and generic ones through their pre-passes over [0, n0)."""
    sr, n, cut = 48000, 40000, 23456
    bank, P = make_bank(20, sr, 123, 0.8, fm=True)
    for v in bank._voices:
        if v.m_wave == 4:
            v.m_pw = 0.1
    force_some_generic(bank, P)
    assert bank.count_generic() > 0 and bank.count_scan() > 0 and bank.count_generic() + bank.count_scan() < 20
    whole = bank.render(n).cpu().numpy()
    tail = bank.render(n - cut, start=cut).cpu().numpy()
    off = np.abs(tail - whole[2 * cut:]) > 1e-6
    assert off.sum() <= 2, "%d samples differ between a whole render and one started at frame %d" % (off.sum(), cut)
    got = dev_pcm16(bank.render(n - cut, start=cut))
    bad = np.abs(got.astype(np.int64) - oracle_bank_pcm(P, n, sr)[2 * cut:]) > 1
    assert bad.mean() < 2e-5


def test_config3_128_voice_chord_bank(cuda):
    """BASELINE config 3: 128-voice chord bank, mixed waveforms, per-voice ADSR, stereo 48 kHz
    (10 s)."""
    sr, seconds = 48000, 10.0
    n = int(sr * seconds)
    bank, P = make_bank(128, sr, 3, seconds, fm=True)
    for v in bank._voices:
        if v.m_wave == 4:
            v.m_pw = 0.1
    s = bank.to_sample(seconds)
    assert (s.nchannels, s.samplerate, len(s)) == (2, sr, n)
    got = np.frombuffer(s.get_frames(), dtype="<i2")
    want = oracle_bank_pcm(P, n, sr)
    bad = np.abs(got.astype(np.int64) - want) > 1
    assert bad.mean() < 2e-5, "%d of %d samples off by more than 1 LSB" % (bad.sum(), bad.size)


def test_config2_fm_saw_adsr_echo_60s_stereo(cuda):
    """BASELINE config 2: Sawtooth carrier FM'd by a Sine, EnvelopeFilter ADSR + EchoFilter,
    60 s stereo 48 kHz.  Mono oscillator chain -> Sample -> stereo()."""
    import synthesizer_b200 as sb
    from oracle import coracle as co
    from oracle.sample_oracle import Sample as OS
    sr, seconds = 48000, 60.0
    n = int(sr * seconds)
    env = (0.5, 1.5, 40.0, 0.6, 10.0)
    lfo = sb.Sine(5.0, 0.03, samplerate=sr)
    car = sb.Sawtooth(220.0, 0.45, fm_lfo=lfo, samplerate=sr)
    chain = sb.EchoFilter(sb.EnvelopeFilter(car, *env), 1.0, 6, 0.35, 0.5)
    s = sb.Sample.from_oscillator(chain, seconds).stereo()
    assert (s.nchannels, len(s)) == (2, n)
    fm = co.osc(0, n, sr, 5.0, 0.03)
    dry = co.envelope(co.osc(3, n, sr, 220.0, 0.45, fm=fm), sr, *env)
    wet = co.echo(dry, sr, 1.0, 6, 0.35, 0.5)
    pcm, bad = co.to_pcm16(wet)
    assert bad == 0
    o = OS.from_raw_frames(pcm.astype("<i2").tobytes(), 2, sr, 1).stereo()
    got = np.frombuffer(s.get_frames(), dtype="<i2").astype(np.int64)
    want = np.frombuffer(o._frames, dtype="<i2").astype(np.int64)
    off = np.abs(got - want) > 1
    # a sawtooth under FM: an edge sample may flip when its phase is within fp64 noise of the wrap
    assert off.sum() <= 4, "%d samples off by more than 1 LSB" % off.sum()


def test_config4_mix_at_4096_resampled_hits(cuda):
    """BASELINE config 4: Sample.mix_at of 4096 resampled drum hits at random offsets into a
    10-minute 48 kHz buffer (mono).  Bit exact against sequential saturating adds."""
    import synthesizer_b200 as sb
    from oracle import coracle as co
    rng = np.random.default_rng(4)
    sr = 48000
    total = 600 * sr
    dst = sb.Sample.from_raw_frames(b"", 2, sr, 1).add_silence(600.0)
    pool, offs, lens, starts, placements = [], [], [], [], []
    pos = 0
    for k in range(4096):
        src_rate = int(rng.choice([22050, 44100, 32000]))
        n = int(rng.integers(2000, 20000))
        t = np.arange(n) / src_rate
        hit = (rng.uniform(8000, 30000) * np.exp(-t * rng.uniform(8, 40)) *
               np.sin(2 * np.pi * rng.uniform(60, 400) * t)).astype(np.int16)
        want_rs = co.pcm16_ratecv(hit, 1, src_rate, sr)
        h = sb.Sample.from_raw_frames(hit.astype("<i2").tobytes(), 2, src_rate, 1).resample(sr)
        if k % 512 == 0:
            np.testing.assert_array_equal(np.frombuffer(h.get_frames(), dtype="<i2"), want_rs)
        when = float(rng.uniform(0, 599.0))
        placements.append((when, h))
        pool.append(want_rs)
        offs.append(pos)
        lens.append(len(want_rs))
        starts.append(int(sr * when))
        pos += len(want_rs)
    dst.mix_at_many(placements)
    want = co.pcm16_mix_at_many(np.zeros(total, dtype=np.int16), np.concatenate(pool), offs, lens, starts)
    got = np.frombuffer(dst.get_frames(), dtype="<i2")
    assert len(got) == total
    np.testing.assert_array_equal(got, want)
    # the same through the batched resample (one launch for all 4096 hits)
    rng = np.random.default_rng(4)
    raw_hits, whens = [], []
    for k in range(4096):
        src_rate = int(rng.choice([22050, 44100, 32000]))
        n = int(rng.integers(2000, 20000))
        t = np.arange(n) / src_rate
        hit = (rng.uniform(8000, 30000) * np.exp(-t * rng.uniform(8, 40)) *
               np.sin(2 * np.pi * rng.uniform(60, 400) * t)).astype(np.int16)
        raw_hits.append(sb.Sample.from_raw_frames(hit.astype("<i2").tobytes(), 2, src_rate, 1))
        whens.append(float(rng.uniform(0, 599.0)))
    sb.Sample.resample_many(raw_hits, sr)
    assert all(h.samplerate == sr for h in raw_hits)
    dst2 = sb.Sample.from_raw_frames(b"", 2, sr, 1).add_silence(600.0)
    dst2.mix_at_many(list(zip(whens, raw_hits)))
    np.testing.assert_array_equal(np.frombuffer(dst2.get_frames(), dtype="<i2"), want)


def test_host_buffer_entry_points(cuda):
    """The host-pointer C-ABI calls (what a non-torch caller binds)."""
    from synthesizer_b200 import _cabi as C
    sr, n = 48000, 24000
    bank, P = make_bank(9, sr, 11, 0.5, waves=[0, 1], fm=False)
    arr = bank.voice_array()
    out = np.empty(2 * n, dtype=np.int16)
    ovf = ctypes.c_ulonglong(0)
    C.call("sb_host_bank_render_pcm16", arr, len(arr), n, float(sr), 32767.0,
           out.ctypes.data_as(ctypes.c_void_p), ctypes.byref(ovf))
    assert ovf.value == 0
    assert_within_lsb(out, oracle_bank_pcm(P, n, sr), 1, "host bank")
    np.testing.assert_array_equal(out, dev_pcm16(bank.render(n)))


def test_host_bank_render_from_two_host_threads(cuda):
    """sb_host_bank_render_pcm16 keeps its workspaces per device behind a mutex: two host threads
    calling it at once (different banks, different sizes, so that the workspaces are regrown under
    each other's feet) each get the bytes they get alone."""
    import threading
    from synthesizer_b200 import _cabi as C
    sr = 48000
    jobs = []
    for seed, nv, n in ((3, 40, 30000), (4, 700, 12000)):
        bank, _ = make_bank(nv, sr, seed, n / sr, fm=True)
        arr = bank.voice_array()
        alone = np.empty(2 * n, dtype=np.int16)
        C.call("sb_host_bank_render_pcm16", arr, len(arr), n, float(sr), 32767.0, alone.ctypes.data_as(ctypes.c_void_p), None)
        jobs.append((arr, n, alone))
    results = {}

    def work(i):
        import torch
        torch.cuda.set_device(0)
        arr, n, _ = jobs[i]
        for rep in range(6):
            out = np.empty(2 * n, dtype=np.int16)
            C.call("sb_host_bank_render_pcm16", arr, len(arr), n, float(sr), 32767.0, out.ctypes.data_as(ctypes.c_void_p), None)
            results[(i, rep)] = out

    threads = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for (i, rep), out in results.items():
        np.testing.assert_array_equal(out, jobs[i][2], err_msg="thread %d, call %d" % (i, rep))
    # the stream variant and the up-front reservation
    import torch
    side = torch.cuda.Stream()
    arr, n, alone = jobs[0]
    C.call("sb_host_reserve", 4096, 48000, float(sr))
    out = np.empty(2 * n, dtype=np.int16)
    C.call("sb_host_bank_render_pcm16_on_stream", arr, len(arr), n, float(sr), 32767.0, out.ctypes.data_as(ctypes.c_void_p),
           None, ctypes.c_void_p(side.cuda_stream))
    np.testing.assert_array_equal(out, alone)


def test_config5_share_full_size_properties(cuda):
    """BASELINE config 5, one GPU's share (8192 FM voices x 10 s @ 48 kHz) at FULL size.
    The serial oracle cannot cover 3.9e9 voice-samples in a test, so: (1) 48 voices drawn from
    the bank are rendered alone at full length and compared with the oracle (+-1 LSB at the
    amplitude they have in the mix, and again normalised to full scale), (2) linearity: the
    bank's mix equals the sum of the mixes of its 8 slices (voice sharding is exact up to fp32
    summation order), (3) the PCM conversion of the full mix reports no overflow."""
    import bench
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C
    from oracle import coracle as co
    sr, frames, nv = bench.SR, int(bench.SR * bench.SECONDS), bench.VOICES_PER_GPU
    p = bench.voice_params(nv)
    arr = bench.build_voices(p, 0, nv, nv)
    bank = sb.VoiceBank.from_voice_array(list(arr), sr)
    assert bank.count_generic() == 0
    whole = bank.render(frames)
    # (2) linearity over 8 voice slices
    acc = None
    for r in range(8):
        b, e = sb.voicebank.shard_range(nv, r, 8)
        part = bank.render(frames, begin=b, end=e)
        acc = part if acc is None else acc + part
    err = float((acc - whole).abs().max())
    assert err < 2e-6, "sum of shard mixes differs from the whole mix by %g" % err
    # (3) conversion
    pcm = dev_pcm16(whole)
    assert np.abs(pcm).max() > 100            # the mix is not silent
    # (1) a sample of voices against the oracle, full length
    rng = np.random.default_rng(99)
    pick = sorted(rng.choice(nv, 48, replace=False).tolist())
    P = np.zeros((len(pick), 18))
    sub = []
    for i, v in enumerate(pick):
        a = arr[v]
        sub.append(a)
        P[i] = [0, a.c_freq, a.c_amp, a.c_phase, 0, 0.1, 0, a.m_freq, a.m_amp, a.m_phase, 0,
                p["attack"][v], p["decay"][v], p["sustain"][v], p["sustain_level"][v], p["release"][v],
                a.gain_l, a.gain_r]
    small = sb.VoiceBank.from_voice_array(sub, sr)
    got = small.render(frames)
    left, right = co.voice_bank(P, frames, sr)
    want = np.empty(2 * frames)
    want[0::2], want[1::2] = left, right
    diff = np.abs(got.cpu().numpy().astype(np.float64) - want)
    # each voice has amplitude ~1e-3/voice in the 8192-voice mix: compare at full scale instead
    scale = 32767.0 / np.abs(want).max()
    assert diff.max() * scale < 1.0, "worst error %.3f LSB (normalised to full scale)" % (diff.max() * scale)
