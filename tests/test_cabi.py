"""
CPU-only checks of the boundary: the library loads, exports every symbol the header declares
(and nothing the binding lacks), the host-only helpers (time base, ADSR boundaries) agree with
the serial loops they stand in for, and argument validation fails loudly.  No GPU calls.
"""
import ctypes
import math
import os
import re

import numpy as np
import pytest

from synthesizer_b200 import _cabi as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "synth_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sb_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    syms = header_symbols()
    assert len(syms) > 40
    for s in syms:
        assert hasattr(C.lib, s), "header declares %s but the library does not export it" % s
    assert sorted(C.EXPORTED) == syms, "ctypes binding and header disagree: %s" % (
        set(C.EXPORTED) ^ set(syms))
    assert C.lib.sb_abi_version() == 2


def test_struct_sizes_match_the_c_side():
    # sizes the kernels assume (sb_voice is copied to shared memory word by word)
    assert ctypes.sizeof(C.TimeSeg) == 24
    assert ctypes.sizeof(C.Envelope) == 80
    assert ctypes.sizeof(C.Voice) == 168
    assert ctypes.sizeof(C.OscParams) == 56


@pytest.mark.parametrize("inc", [1 / 48000, 2 * math.pi / 44100, 1 / 22050, 2 * math.pi / 96000, 0.1])
def test_timebase_is_bit_identical_to_serial_accumulation(inc):
    segs = (C.TimeSeg * 256)()
    n = C.lib.sb_timebase_build(inc, 1 << 40, segs, 256)
    assert 0 < n < 100
    N = 3_000_000
    serial = np.concatenate([[0.0], np.cumsum(np.full(N - 1, inc))])      # cumsum adds sequentially
    starts = np.array([segs[i].n_start for i in range(n)])
    ts = np.array([segs[i].t_start for i in range(n)])
    st = np.array([segs[i].step for i in range(n)])
    idx = np.searchsorted(starts, np.arange(N), side="right") - 1
    closed = ts[idx] + (np.arange(N) - starts[idx]).astype(np.float64) * st[idx]
    assert np.array_equal(serial, closed)
    for k in (0, 1, 2, 12345, N - 1):
        assert C.lib.sb_timebase_at(segs, n, k) == serial[k]
    lim = serial[54321]
    assert C.lib.sb_timebase_first_ge(segs, n, 0, lim) == 54321
    assert C.lib.sb_timebase_first_ge(segs, n, 0, np.nextafter(lim, 1e9)) == 54322
    assert C.lib.sb_timebase_first_ge(segs, n, 60000, lim) == 60000


@pytest.mark.parametrize("t0,inc", [(0.3 * 2 * math.pi, 2 * math.pi * 441.0 / 44100), (0.77 / 50.0, 1 / 44100),
                                    (123.456, 1 / 48000), (1e-9, 1 / 96000)])
def test_timebase_from_a_start_value(t0, inc):
    segs = (C.TimeSeg * 256)()
    n = C.lib.sb_timebase_build_from(t0, inc, 1 << 40, segs, 256)
    assert 0 < n < 120
    N = 2_000_000
    serial = np.empty(N)
    serial[0] = t0
    acc = np.cumsum(np.concatenate([[t0], np.full(N - 1, inc)]))      # sequential adds starting from t0
    serial[:] = acc
    starts = np.array([segs[i].n_start for i in range(n)])
    ts = np.array([segs[i].t_start for i in range(n)])
    st = np.array([segs[i].step for i in range(n)])
    idx = np.searchsorted(starts, np.arange(N), side="right") - 1
    closed = ts[idx] + (np.arange(N) - starts[idx]).astype(np.float64) * st[idx]
    assert np.array_equal(serial, closed)
    assert C.lib.sb_timebase_build_from(-1.0, inc, 10, segs, 256) < 0


@pytest.mark.parametrize("adsr", [(0.05, 0.1, 0.3, 0.6, 0.2), (0.0, 0.1, 0.0, 0.5, 0.1), (0.1, 0.0, 0.2, 1.0, 0.0),
                                  (0.013, 0.027, 0.111, 0.35, 0.3333), (1 / 48000, 2 / 48000, 0.0, 0.0, 0.01)])
def test_envelope_boundaries_match_the_serial_generator(adsr):
    """Replays upstream's EnvelopeFilter loop (oracle) and compares segment boundaries."""
    from oracle import coracle as co
    sr = 48000
    env = C.Envelope()
    C.call("sb_envelope_build", float(sr), *[float(x) for x in adsr], ctypes.byref(env))
    n = int(sum(adsr[:3] + adsr[4:]) * sr) + 100
    ones = np.ones(n)
    shape = co.envelope(ones, sr, *adsr)                      # the envelope itself
    cut = co.envelope(ones, sr, *adsr, stop_at_end=True)
    assert len(cut) == env.n_r + env.extra
    k = np.arange(n)
    model = np.where(k < env.n_a, k * env.a_slope,
             np.where(k < env.n_d, 1.0 + (k - env.n_a) * env.d_slope,
              np.where(k < env.n_s, env.sustain_level,
               np.where(k < env.n_r, env.sustain_level + (k - env.n_s) * env.r_slope,
                np.where((k == env.n_r) & (env.extra == 1), env.amp_extra, 0.0)))))
    np.testing.assert_allclose(model, shape, rtol=0, atol=1e-9)
    # the zero / non-zero pattern (where segments start and stop) is exact
    assert np.array_equal(model != 0, shape != 0)


def test_release_replay_equals_the_serial_accumulation():
    """sb_envelope_build decides the "one more sample" of the release from the amplitude that
    upstream's loop has accumulated by then (amp += step, once per sample).  The library jumps
    through that accumulation one binade at a time; here it is replayed addition by addition."""
    rng = np.random.default_rng(8)
    cases = [(0.05, 0.1, 0.3, 0.6, 0.2), (0.0, 0.0, 0.0, 1.0, 3.0), (0.01, 0.02, 0.0, 1e-9, 0.5), (0.2, 0.1, 0.1, 0.333, 1e-4)]
    cases += [tuple(float(x) for x in (rng.uniform(0, 0.3), rng.uniform(0, 0.3), rng.uniform(0, 0.5),
                                        rng.uniform(0, 1), rng.uniform(1e-3, 2.5))) for _ in range(24)]
    for sr in (44100.0, 48000.0):
        for adsr in cases:
            env = C.Envelope()
            C.call("sb_envelope_build", sr, *adsr, ctypes.byref(env))
            amp, step = adsr[3], env.r_slope
            for _ in range(env.n_r - env.n_s):
                amp = amp + step
            assert env.extra == (1 if amp > 0.0 else 0), (sr, adsr, amp)
            if env.extra:
                assert env.amp_extra == amp, (sr, adsr, amp, env.amp_extra)


def test_bad_arguments_fail_loudly():
    segs = (C.TimeSeg * 4)()
    assert C.lib.sb_timebase_build(-1.0, 10, segs, 4) < 0
    assert b"bad arguments" in C.lib.sb_last_error()
    with pytest.raises(C.SbError):
        C.call("sb_envelope_build", 48000.0, -1.0, 0.0, 0.0, 0.5, 0.0, ctypes.byref(C.Envelope()))
    plan = C.BankPlan()
    with pytest.raises(C.SbError):
        C.call("sb_bank_plan_make", 0, 100, 0, 148, 40, ctypes.byref(plan))
    assert C.lib.sb_pcm16_ratecv_frames(8, 44100, 48000) == 8
    assert C.lib.sb_pcm16_ratecv_frames(8, 48000, 22050) == 4
    assert C.lib.sb_pcm16_ratecv_frames(0, 48000, 22050) == 0


def test_bank_plan_covers_the_work():
    for nv, frames in [(1, 100), (1, 2_880_000), (128, 480_000), (8192, 480_000), (65536, 480_000), (3, 257)]:
        plan = C.BankPlan()
        C.call("sb_bank_plan_make", nv, frames, 0, 148, 40, ctypes.byref(plan))
        assert plan.nprefix == 0 and plan.n0 == 0
        shifted = C.BankPlan()
        C.call("sb_bank_plan_make", nv, frames, 5 * plan.span_len + 1, 148, 40, ctypes.byref(shifted))
        assert shifted.nprefix == 6 and shifted.carry_bytes >= plan.carry_bytes
        assert shifted.off_fastv >= nv * (6 + shifted.nspans) * 8
        assert plan.voices_per_group * plan.ngroups >= nv
        assert plan.span_len % 256 == 0 and plan.span_len * plan.nspans >= frames
        assert plan.span_len * (plan.nspans - 1) < frames
        assert plan.fast_span_len % 256 == 0 and 1 <= plan.fast_span_len // 256 <= 8
        assert plan.fast_span_len * plan.fast_nspans >= frames > plan.fast_span_len * (plan.fast_nspans - 1)
        assert plan.off_fastv >= nv * plan.nspans * 8 and plan.off_tab >= plan.off_fastv + nv * 128
        assert plan.carry_bytes >= plan.off_tab + nv * 40 * 16
        assert 1 <= plan.voices_per_group <= 64


def test_bank_plan_hint_sizes_the_scratch():
    """Without a hint the scratch is sized as if every voice needed the modulator tile table;
    the hint shrinks it to the class-2 voices and records which kernels can be skipped."""
    nv, frames = 1000, 480_000
    plan = C.BankPlan()
    C.call("sb_bank_plan_make", nv, frames, 0, 148, 40, ctypes.byref(plan))
    assert (plan.generic_voices, plan.edged_voices, plan.scan_voices) == (-1, -1, -1)
    worst = plan.carry_bytes
    rows = (frames + 511) // 512
    assert worst >= nv * rows * 8
    C.call("sb_bank_plan_hint", ctypes.byref(plan), 0, nv, 0, 0)
    none = plan.carry_bytes
    assert none <= worst - nv * rows * 8 and (plan.generic_voices, plan.edged_voices, plan.scan_voices) == (0, 0, 0)
    C.call("sb_bank_plan_hint", ctypes.byref(plan), 3, -1, 10, 7)
    assert none + 7 * rows * 8 <= plan.carry_bytes <= none + 7 * rows * 8 + 256
    C.call("sb_bank_plan_hint", ctypes.byref(plan), -1, -1, -1, -1)
    assert plan.carry_bytes == worst
    with pytest.raises(C.SbError):
        C.call("sb_bank_plan_hint", ctypes.byref(plan), 0, 0, 0, nv + 1)
    # a start offset adds the tiles over [0, n0)
    shifted = C.BankPlan()
    C.call("sb_bank_plan_make", nv, frames, 512 * 10 + 1, 148, 40, ctypes.byref(shifted))
    C.call("sb_bank_plan_hint", ctypes.byref(shifted), 0, -1, 0, 7)
    C.call("sb_bank_plan_hint", ctypes.byref(plan), 0, -1, 0, 7)
    assert shifted.carry_bytes - plan.carry_bytes >= 7 * 11 * 8 - 256


def test_voice_classes_on_the_host():
    """sb_voice_is_fast: 1 = closed-form fast path, 2 = tile-scan fast path, 0 = generic."""
    v = C.Voice()
    v.c_wave, v.c_freq, v.c_amp, v.c_pw = C.WAVE_SINE, 440.0, 0.5, 0.1
    v.m_wave = C.WAVE_NONE
    cls = lambda: C.lib.sb_voice_is_fast(ctypes.byref(v), 48000.0)
    assert cls() == 1
    v.m_wave, v.m_freq, v.m_amp, v.m_pw = C.WAVE_SINE, 5.0, 0.05, 0.1
    assert cls() == 1
    for w in (C.WAVE_TRIANGLE, C.WAVE_SQUARE, C.WAVE_SAWTOOTH, C.WAVE_PULSE):
        v.m_wave = w
        assert cls() == 2
    v.m_phase = -0.25
    assert cls() == 0                       # negative modulator phase: periodic rules do not hold
    v.m_phase, v.m_amp = 0.0, 30.0          # 2*pi*440/48000 * 30 = 1.7 rad per unit of modulator output:
    assert cls() == 2                       # strong FM, still fast for a Sine / Triangle carrier
    v.c_wave = C.WAVE_SAWTOOTH
    assert cls() == 0                       # ... generic for an edged carrier (limit 1.5 rad)
    v.c_wave, v.c_freq = C.WAVE_SINE, 1760.0
    assert cls() == 0                       # 6.9 rad: beyond the strong-FM limit of 6
    v.c_freq = 440.0
    v.m_amp, v.c_wave, v.c_phase = 0.05, C.WAVE_SQUARE, -0.1
    assert cls() == 0                       # edged carrier starting at a negative phase
    v.c_phase = 0.3
    assert cls() == 2
    v.m_wave = C.WAVE_SINE
    assert cls() == 1


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import synthesizer_b200 as sb
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sb.Sine(440).render(16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        sb.Sample.from_raw_frames(b"\0\0", 2, 44100, 1)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "synthesizer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("no CPU", ""), "%s mentions the oracle" % f
