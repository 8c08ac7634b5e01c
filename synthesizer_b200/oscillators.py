"""
Oscillators and filters with the API of ``synthplayer.oscillators``, rendered on a B200.

Drop-in surface (names, argument order and meaning follow the published synthplayer package;
the source is absent from /root/reference, so no file:line can be cited — see DESIGN.md):
``Oscillator.blocks()`` yields lists of ``params.norm_osc_blocksize`` floats, ``generator()``
single floats; filters wrap oscillators by construction.  What differs is underneath: an
oscillator is a node of a render graph, and a request for samples [n0, n0+count) is answered
by CUDA kernels through the C ABI (include/synth_b200.h) — per-sample-index closed forms on
the exact time base instead of serial Python loops.  ``render(nframes)`` returns the samples
as a float32 CUDA tensor without the host round trip that blocks()/generator() imply.

There is no CPU implementation behind these classes.
"""
import ctypes
import logging
import itertools
import math
import random
import sys
from math import log

import torch

from . import _cabi as C
from . import _device as D
from . import params

__all__ = ["Oscillator", "Filter", "Sine", "Triangle", "Square", "SquareH", "Harmonics", "Pulse", "Sawtooth",
           "SawtoothH", "WhiteNoise", "Linear", "FastSine", "FastPulse", "FastTriangle", "FastSawtooth",
           "FastSquare", "Semicircle", "Pointy", "EnvelopeFilter", "MixingFilter", "AmpModulationFilter", "DelayFilter",
           "EchoFilter", "ClipFilter", "AbsFilter", "NullFilter"]

_FM_SEEK_CHUNK = 1 << 22


class Oscillator:
    """Base class (synthplayer.oscillators.Oscillator)."""

    def __init__(self, source=None, samplerate=None):
        self._samplerate = samplerate or source._samplerate
        self._source = source
        self._cache_n0 = 0
        self._cache = None            # float32 CUDA tensor holding samples [_cache_n0, ...)
        self._cache64_n0 = 0
        self._cache64 = None          # the same for the fp64 values handed to an FM consumer (_fetch64)

    # ---- upstream API ---------------------------------------------------------------
    def __iter__(self):
        return self.blocks()

    def blocks(self):
        """Lists of params.norm_osc_blocksize floats, forever (or until a filter ends)."""
        bs = params.norm_osc_blocksize
        pos = 0
        chunk = params.render_chunk_min
        while True:
            total = self._length()
            n = chunk if total is None else min(chunk, total - pos)
            if n <= 0:
                return
            data = self._fetch(pos, pos + n).cpu().tolist()
            for i in range(0, n, bs):
                yield data[i:i + bs]
            pos += n
            chunk = min(chunk * 2, params.render_chunk_max)

    def generator(self):
        """Single float samples (the older synthplayer interface)."""
        return itertools.chain.from_iterable(self.blocks())

    # ---- backend API ------------------------------------------------------------------
    def render(self, nframes, start=0):
        """Samples [start, start+nframes) as a float32 CUDA tensor (shorter if the stream ends)."""
        D.require_cuda()
        total = self._length()
        end = start + int(nframes)
        if total is not None:
            end = min(end, total)
        if end <= start:
            return D.empty(0, torch.float32)
        return self._fetch(start, end)

    def _length(self):
        """Total number of samples of a finite stream, or None."""
        return None

    def _fetch(self, a, b):
        """Samples [a, b), reusing the overlap with the previous request so that sequential
        (and look-back) consumers never re-render and stateful nodes advance monotonically."""
        c = self._cache
        if c is not None and self._cache_n0 <= a <= self._cache_n0 + c.numel():
            c_end = self._cache_n0 + c.numel()
            if b <= c_end:
                return c[a - self._cache_n0:b - self._cache_n0]
            new = self._render_range(c_end, b - c_end)
            out = torch.cat((c[a - self._cache_n0:], new)) if a < c_end else new
        else:
            out = self._render_range(a, b - a)
        self._cache_n0, self._cache = a, out
        return out

    def _render_range(self, n0, count):
        raise NotImplementedError

    def _fetch64(self, a, b):
        """Samples [a, b) in fp64: this node as another oscillator's FM modulator.  The modulator's
        values are summed over the whole render, and float32 rounding of a slow modulator does
        not average out there (measured: 200 - 1200 LSB of carrier error after 60 s), so the nodes
        an LFO is usually made of - waveform oscillators, Linear, and the envelope / mixing /
        amplitude-modulation / clip / abs / delay filters over them - compute in fp64 as upstream's
        Python floats do.  Any other node hands over its float32 values."""
        c = self._cache64
        if c is not None and self._cache64_n0 <= a <= self._cache64_n0 + c.numel():
            # the overlap with the previous request is reused (a modulator shared by several carriers is
            # rendered once; a stateful one advances monotonically), as _fetch does for float32
            c_end = self._cache64_n0 + c.numel()
            if b <= c_end:
                return c[a - self._cache64_n0:b - self._cache64_n0]
            new = self._values64(c_end, b)
            out = torch.cat((c[a - self._cache64_n0:], new)) if a < c_end else new
        else:
            out = self._values64(a, b)
        self._cache64_n0, self._cache64 = a, out
        return out

    def _values64(self, a, b):
        """Samples [a, b) in fp64, rendered now (nodes that compute in fp64 override this)."""
        return self._fetch(a, b).double()


def _osc_params(wave, frequency, amplitude, phase, bias, pulsewidth, samplerate):
    p = C.OscParams()
    p.wave, p.frequency, p.amplitude, p.phase = wave, frequency, amplitude, phase
    p.bias, p.pulsewidth, p.samplerate = bias, pulsewidth, samplerate
    return p


class _WaveOscillator(Oscillator):
    """Sine / Triangle / Square / Sawtooth / Pulse with optional FM (and PWM for Pulse)."""

    _wave = None

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        super().__init__(samplerate=samplerate or params.norm_samplerate)
        self.frequency = frequency
        self.amplitude = amplitude
        self.bias = bias
        self._phase = phase
        self._fm_lfo = fm_lfo
        self._pwm_lfo = None
        self.pulsewidth = 0.1
        self._fm_pos = 0
        self._fm_carry = None

    def _timebase(self):
        return D.timebase_rad(self._samplerate) if self._wave in (C.WAVE_SINE, C.WAVE_POINTY) else D.timebase_cyc(self._samplerate)

    def _length(self):
        lens = [x._length() for x in (self._fm_lfo, self._pwm_lfo) if x is not None]
        lens = [x for x in lens if x is not None]
        return min(lens) if lens else None

    def _fm_prefix(self, n0, count, tb, out=True):
        """Exclusive prefix sum of lfo*step over [n0, n0+count) continuing the running carry."""
        fmsum = D.empty(count, torch.float64)
        scratch = D.empty(C.lib.sb_fm_prefix_scratch_bytes(count), torch.uint8)
        # the modulator's values in fp64 (see Oscillator._fetch64; the plain waveform modulators never
        # get here: they go through the fused kernel, _as_single_voice)
        lfo = self._fm_lfo._fetch64(n0, n0 + count)
        if lfo.numel() < count:                       # a finite modulator that has ended: silence from there on
            lfo = torch.cat((lfo, D.zeros(count - lfo.numel(), torch.float64)))
        C.call("sb_fm_prefix_f64_from_f64", D.ptr(lfo.contiguous()), D.ptr(tb.dev()), tb.nsegs, n0, count,
               D.ptr(self._fm_carry), D.ptr(fmsum), D.ptr(scratch), D.stream())
        self._fm_pos = n0 + count
        return fmsum

    def _fm_seek(self, n0, tb):
        """Bring the running FM sum to sample n0 (restarting from 0 when going backwards)."""
        if self._fm_carry is None or n0 < self._fm_pos:
            self._fm_carry = D.zeros(1, torch.float64)
            self._fm_pos = 0
        while self._fm_pos < n0:
            self._fm_prefix(self._fm_pos, min(_FM_SEEK_CHUNK, n0 - self._fm_pos), tb)

    def _as_single_voice(self):
        """FM by a plain waveform oscillator is one voice of the fused bank kernel: the modulator
        is then evaluated in fp64 (or in closed form) inside that kernel instead of passing
        through a float32 node buffer, whose rounding does not average out over a slow LFO."""
        lfo = self._fm_lfo
        if lfo is None or self._pwm_lfo is not None or type(lfo)._render_range is not _WaveOscillator._render_range:
            return None
        from .voicebank import VoiceBank
        try:
            return VoiceBank(self._samplerate).add(self, 1.0, 0.0)
        except ValueError:
            return None

    def _values64(self, a, b):
        if type(self)._render_range is not _WaveOscillator._render_range:      # Harmonics: float32 values
            return super()._values64(a, b)
        return self._render_range64(a, b - a)

    def _render_range64(self, n0, count):
        """Samples [n0, n0+count) in fp64: this oscillator as another one's FM modulator."""
        tb = self._timebase()
        fmsum = None
        if self._fm_lfo is not None:
            self._fm_seek(n0, tb)
            fmsum = self._fm_prefix(n0, count, tb)
        pwm = self._pwm_lfo._fetch(n0, n0 + count) if self._pwm_lfo is not None else None
        out = D.empty(count, torch.float64)
        p = _osc_params(self._wave, self.frequency, self.amplitude, self._phase, self.bias,
                        self.pulsewidth, self._samplerate)
        C.call("sb_osc_render_f64", ctypes.byref(p), D.ptr(tb.dev()), tb.nsegs, n0, count,
               D.ptr(fmsum), D.ptr(pwm), D.ptr(out), D.stream())
        return out

    def _render_range(self, n0, count):
        bank = self._as_single_voice()
        if bank is not None:
            mix = bank.render(count, start=n0)
            out = D.empty(count, torch.float32)
            C.call("sb_channel_f32", D.ptr(mix), 2, 0, count, D.ptr(out), D.stream())
            return out
        tb = self._timebase()
        fmsum = None
        if self._fm_lfo is not None:
            self._fm_seek(n0, tb)
            fmsum = self._fm_prefix(n0, count, tb)
        pwm = self._pwm_lfo._fetch(n0, n0 + count) if self._pwm_lfo is not None else None
        out = D.empty(count, torch.float32)
        p = _osc_params(self._wave, self.frequency, self.amplitude, self._phase, self.bias,
                        self.pulsewidth, self._samplerate)
        C.call("sb_osc_render_f32", ctypes.byref(p), D.ptr(tb.dev()), tb.nsegs, n0, count,
               D.ptr(fmsum), D.ptr(pwm), D.ptr(out), D.stream())
        return out


class Sine(_WaveOscillator):
    """Sine wave oscillator (synthplayer.oscillators.Sine)."""
    _wave = C.WAVE_SINE


class Triangle(_WaveOscillator):
    """Perfect triangle wave oscillator (synthplayer.oscillators.Triangle)."""
    _wave = C.WAVE_TRIANGLE


class Square(_WaveOscillator):
    """Perfect square wave oscillator (synthplayer.oscillators.Square)."""
    _wave = C.WAVE_SQUARE


class Sawtooth(_WaveOscillator):
    """Perfect sawtooth wave oscillator (synthplayer.oscillators.Sawtooth)."""
    _wave = C.WAVE_SAWTOOTH


class Pulse(_WaveOscillator):
    """Pulse (narrow square) oscillator with optional pulse width modulation
    (synthplayer.oscillators.Pulse)."""
    _wave = C.WAVE_PULSE

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, pulsewidth=0.1,
                 fm_lfo=None, pwm_lfo=None, samplerate=0):
        assert 0 <= pulsewidth <= 1
        super().__init__(frequency, amplitude, phase, bias, fm_lfo, samplerate)
        self.pulsewidth = pulsewidth
        self._pwm_lfo = pwm_lfo


class Semicircle(_WaveOscillator):
    """Semicircle half ('W3'), with optional FM (synthplayer.oscillators.Semicircle; formula
    recalled with low confidence: sqrt(0.25 - ft*ft) * 2*amplitude + bias, ft = frac(phase) - 0.5)."""
    _wave = C.WAVE_SEMICIRCLE


class Pointy(_WaveOscillator):
    """Pointy wave ('inverted cosine', 'W2'), with optional FM (synthplayer.oscillators.Pointy;
    formula recalled with low confidence: vv = 1 - |cos x|, copysign(vv*vv, sin x)*amplitude + bias)."""
    _wave = C.WAVE_POINTY


class Harmonics(_WaveOscillator):
    """Sum of sine harmonics [(n, amplitude_factor), ...] below Nyquist, with optional FM
    (synthplayer.oscillators.Harmonics)."""
    _wave = C.WAVE_SINE

    def __init__(self, frequency, harmonics, amplitude=1.0, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        super().__init__(frequency, amplitude, phase, bias, fm_lfo, samplerate)
        nyquist = self._samplerate / 2.0
        self.harmonics = [(n, a) for n, a in harmonics if n * frequency <= nyquist]
        if len(self.harmonics) > 128:
            raise ValueError("at most 128 harmonics are supported")

    def _render_range(self, n0, count):
        tb = self._timebase()
        fmsum = None
        if self._fm_lfo is not None:
            self._fm_seek(n0, tb)
            fmsum = self._fm_prefix(n0, count, tb)
        out = D.empty(count, torch.float32)
        p = _osc_params(C.WAVE_SINE, self.frequency, self.amplitude, self._phase, self.bias, 0.0, self._samplerate)
        nh = len(self.harmonics)
        hn = (ctypes.c_int32 * max(1, nh))(*[int(n) for n, _ in self.harmonics])
        ha = (ctypes.c_float * max(1, nh))(*[float(a) for _, a in self.harmonics])
        C.call("sb_harmonics_render_f32", ctypes.byref(p), hn, ha, nh, D.ptr(tb.dev()), tb.nsegs, n0, count,
               D.ptr(fmsum), D.ptr(out), D.stream())
        return out


class SquareH(Harmonics):
    """Square wave from odd sine harmonics (synthplayer.oscillators.SquareH)."""

    def __init__(self, frequency, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        harmonics = [(n, 1.0 / n) for n in range(1, num_harmonics * 2, 2)]
        super().__init__(frequency, harmonics, amplitude, phase, bias, fm_lfo, samplerate)


class SawtoothH(Harmonics):
    """Sawtooth wave from sine harmonics (synthplayer.oscillators.SawtoothH)."""

    def __init__(self, frequency, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        harmonics = [(n, 1.0 / n) for n in range(1, num_harmonics + 1)]
        super().__init__(frequency, harmonics, -amplitude, phase + 0.5, bias, fm_lfo, samplerate)


class _FastOscillator(Oscillator):
    """The Fast* oscillators of upstream: no FM, t starts at a phase-dependent value and is
    advanced serially; here evaluated per sample index on a time base built for that start
    value, with frequency folded the way each upstream class folds it."""
    _wave = None

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, samplerate=0):
        super().__init__(samplerate=samplerate or params.norm_samplerate)
        self._frequency = frequency
        self._phase = phase
        self.amplitude = amplitude
        self.bias = bias
        self._pulsewidth = 0.1
        if phase < 0:
            raise ValueError("negative phase is not supported for the Fast* oscillators on this backend")

    def _render_range(self, n0, count):
        return self._render(n0, count, torch.float32, "sb_osc_render_f32")

    def _values64(self, a, b):
        # (a Fast* oscillator is a natural LFO: its values in fp64 when it modulates a frequency)
        return self._render(a, b - a, torch.float64, "sb_osc_render_f64")

    def _render(self, n0, count, dtype, entry):
        if self._wave == C.WAVE_SINE:
            # FastSine: t0 = phase*2*pi, increment = 2*pi/(samplerate/frequency), value = sin(t)
            rate = self._samplerate / self._frequency
            tb = D.timebase(2.0 * math.pi / rate, self._phase * 2.0 * math.pi)
            freq = 1.0
        else:
            # FastTriangle & co: t0 = phase/frequency, increment = 1/samplerate, tt = t*frequency
            tb = D.timebase(1.0 / self._samplerate, self._phase / self._frequency)
            freq = self._frequency
        out = D.empty(count, dtype)
        p = _osc_params(self._wave, freq, self.amplitude, 0.0, self.bias, self._pulsewidth, self._samplerate)
        C.call(entry, ctypes.byref(p), D.ptr(tb.dev()), tb.nsegs, n0, count,
               D.ptr(None), D.ptr(None), D.ptr(out), D.stream())
        return out


class FastSine(_FastOscillator):
    """Fast sine wave oscillator, no FM (synthplayer.oscillators.FastSine)."""
    _wave = C.WAVE_SINE


class FastTriangle(_FastOscillator):
    """Fast triangle wave oscillator, no FM (synthplayer.oscillators.FastTriangle)."""
    _wave = C.WAVE_TRIANGLE


class FastSquare(_FastOscillator):
    """Fast square wave oscillator, no FM (synthplayer.oscillators.FastSquare)."""
    _wave = C.WAVE_SQUARE


class FastSawtooth(_FastOscillator):
    """Fast sawtooth wave oscillator, no FM (synthplayer.oscillators.FastSawtooth)."""
    _wave = C.WAVE_SAWTOOTH


class FastPulse(_FastOscillator):
    """Fast pulse wave oscillator, fixed pulse width, no FM/PWM (synthplayer.oscillators.FastPulse)."""
    _wave = C.WAVE_PULSE

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, pulsewidth=0.1, samplerate=0):
        assert 0 <= pulsewidth <= 1
        super().__init__(frequency, amplitude, phase, bias, samplerate)
        self._pulsewidth = pulsewidth


class WhiteNoise(Oscillator):
    """White noise: one uniform draw in [-amplitude, amplitude] held for
    int(samplerate/frequency) samples (synthplayer.oscillators.WhiteNoise).

    Upstream draws from Python's ``random``; here the draws are counter-based (Philox) with a
    seed taken from ``random`` at construction, so ``random.seed()`` still makes a run
    reproducible.  ``noise_source`` (a sequence of uniform draws in [0,1)) injects the draws
    instead — that is how parity tests feed both implementations the same numbers."""

    def __init__(self, frequency, amplitude=1.0, bias=0.0, samplerate=0, noise_source=None):
        super().__init__(samplerate=samplerate or params.norm_samplerate)
        self.frequency = frequency
        self.amplitude = amplitude
        self.bias = bias
        self._cycles = int(self._samplerate / self.frequency)
        if self._cycles < 1:
            raise ValueError("whitenoise frequency cannot be bigger than the sample rate")
        self._seed = random.getrandbits(64)
        self._injected = None
        if noise_source is not None:
            self._injected = torch.as_tensor(list(noise_source) if not hasattr(noise_source, "__array__")
                                             else noise_source, dtype=torch.float32)

    def _length(self):
        return None if self._injected is None else self._injected.numel() * self._cycles

    def _render_range(self, n0, count):
        d0 = n0 // self._cycles
        d1 = (n0 + count - 1) // self._cycles + 1
        if self._injected is not None:
            u = self._injected[d0:d1].to(D.device())
        else:
            u = D.empty(d1 - d0, torch.float32)
            C.call("sb_uniform_f32", self._seed, d0, d1 - d0, D.ptr(u), D.stream())
        out = D.empty(count, torch.float32)
        C.call("sb_noise_render_f32", D.ptr(u), d0, self._cycles, self.amplitude, self.bias,
               n0, count, D.ptr(out), D.stream())
        return out


class Linear(Oscillator):
    """Linear ramp, clamped to [min_value, max_value] (synthplayer.oscillators.Linear)."""

    def __init__(self, startlevel, increment=0.0, min_value=-1.0, max_value=1.0, samplerate=0):
        super().__init__(samplerate=samplerate or params.norm_samplerate)
        self.value = startlevel
        self.increment = increment
        self.min_value = min_value
        self.max_value = max_value

    def _render_range(self, n0, count):
        out = D.empty(count, torch.float32)
        C.call("sb_linear_render_f32", self.value, self.increment, self.min_value, self.max_value,
               n0, count, D.ptr(out), D.stream())
        return out

    def _values64(self, a, b):
        out = D.empty(b - a, torch.float64)
        C.call("sb_linear_render_f64", self.value, self.increment, self.min_value, self.max_value,
               a, b - a, D.ptr(out), D.stream())
        return out


# ------------------------------------------------------------------------------ filters

def _min_len(*lens):
    lens = [x for x in lens if x is not None]
    return min(lens) if lens else None


class Filter(Oscillator):
    """Base of everything that wraps other oscillators (synthplayer.oscillators.Filter)."""

    def __init__(self, sources):
        self._sources = list(sources)
        super().__init__(self._sources[0])

    def _length(self):
        return _min_len(*[s._length() for s in self._sources])


class EnvelopeFilter(Filter):
    """ADSR volume envelope; A, D, S, R in seconds, sustain_level a factor
    (synthplayer.oscillators.EnvelopeFilter).  After the release the output is 0.0 forever,
    or the stream ends there when stop_at_end is set."""

    def __init__(self, source, attack, decay, sustain, sustain_level, release, stop_at_end=False):
        assert attack >= 0 and decay >= 0 and sustain >= 0 and release >= 0
        assert 0 <= sustain_level <= 1
        super().__init__([source])
        self._attack = attack
        self._decay = decay
        self._sustain = sustain
        self._sustain_level = sustain_level
        self._release = release
        self._stop_at_end = stop_at_end
        self._env = D.build_envelope(self._samplerate, attack, decay, sustain, sustain_level, release)

    def _length(self):
        src = self._sources[0]._length()
        if self._stop_at_end:
            return _min_len(src, self._env.n_r + self._env.extra)
        return src

    def _render_range(self, n0, count):
        # beyond the release nothing is pulled from the source (as upstream), only zeros
        end_env = self._env.n_r + self._env.extra
        live = max(0, min(count, end_env - n0))
        out = D.zeros(count, torch.float32) if live < count else D.empty(count, torch.float32)
        if live:
            src = self._sources[0]._fetch(n0, n0 + live)
            C.call("sb_envelope_apply_f32", D.ptr(src), ctypes.byref(self._env), n0, live,
                   D.ptr(out), D.stream())
        return out

    def _values64(self, a, b):
        count = b - a
        end_env = self._env.n_r + self._env.extra
        live = max(0, min(count, end_env - a))
        out = D.zeros(count, torch.float64)
        if live:
            src = self._sources[0]._fetch64(a, a + live).contiguous()
            C.call("sb_envelope_apply_f64", D.ptr(src), ctypes.byref(self._env), a, src.numel(),
                   D.ptr(out), D.stream())
        return out


class MixingFilter(Filter):
    """Sum of all sources (synthplayer.oscillators.MixingFilter)."""

    def __init__(self, *sources):
        super().__init__(sources)

    def _as_bank(self):
        """The sources that are plain voices — [EnvelopeFilter(] waveform oscillator with an
        optional plain waveform FM LFO [)] — go into ONE launch of the fused voice-bank kernel
        instead of one kernel chain each; returns (bank or None, the other sources).  A source
        the bank does not take (PWM, a stop_at_end envelope, a composite LFO, another filter)
        is rendered node by node and added to the bank's mix; the split is logged once per
        filter at INFO level.  Rebuilt per request because upstream oscillators may have their
        public attributes changed between renders."""
        from .voicebank import VoiceBank
        bank, rest = VoiceBank(self._samplerate), []
        for s in self._sources:
            try:
                bank.add(s, 1.0, 0.0)
            except ValueError:
                rest.append(s)
        if len(bank) < 2:
            return None, self._sources
        if rest and not getattr(self, "_split_logged", False):
            self._split_logged = True
            logging.getLogger("synthesizer_b200").info(
                "MixingFilter: %d of %d sources are not plain bank voices and render node by node (%s)",
                len(rest), len(self._sources), ", ".join(sorted({type(s).__name__ for s in rest})))
        return bank, rest

    def _render_range(self, n0, count):
        bank, rest = self._as_bank()
        bufs = []
        if bank is not None:
            mix = bank.render(count, start=n0)
            out = D.empty(count, torch.float32)
            C.call("sb_channel_f32", D.ptr(mix), 2, 0, count, D.ptr(out), D.stream())   # left channel = the mono mix
            if not rest:
                return out
            bufs.append(out)
        bufs += [s._fetch(n0, n0 + count) for s in rest]
        ptrs = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64).to(D.device())
        out = D.empty(count, torch.float32)
        C.call("sb_mix_f32", D.ptr(ptrs), len(bufs), count, D.ptr(out), D.stream())
        return out

    def _values64(self, a, b):
        # (summed in source order, as upstream's sum(values) does)
        parts = [s._fetch64(a, b) for s in self._sources]
        n = min(p.numel() for p in parts)
        out = parts[0][:n].clone()
        for p in parts[1:]:
            out += p[:n]
        return out


class AmpModulationFilter(Filter):
    """Amplitude modulation by another oscillator (synthplayer.oscillators.AmpModulationFilter)."""

    def __init__(self, source, modulator):
        super().__init__([source, modulator])

    def _render_range(self, n0, count):
        a = self._sources[0]._fetch(n0, n0 + count)
        b = self._sources[1]._fetch(n0, n0 + count)
        out = D.empty(count, torch.float32)
        C.call("sb_mul_f32", D.ptr(a), D.ptr(b), count, D.ptr(out), D.stream())
        return out

    def _values64(self, a, b):
        x, y = self._sources[0]._fetch64(a, b), self._sources[1]._fetch64(a, b)
        n = min(x.numel(), y.numel())
        return x[:n] * y[:n]


class DelayFilter(Filter):
    """Delays the source by int(samplerate*seconds) samples of silence, or skips that many
    samples for a negative delay (synthplayer.oscillators.DelayFilter)."""

    def __init__(self, source, seconds):
        super().__init__([source])
        self._seconds = seconds
        if seconds < 0.0:
            self._shift = -int(-self._samplerate * seconds)
        else:
            self._shift = int(self._samplerate * seconds)

    def _length(self):
        src = self._sources[0]._length()
        return None if src is None else max(0, src + self._shift)

    def _render_range(self, n0, count):
        a, b = n0 - self._shift, n0 + count - self._shift       # source indices
        if b <= 0:
            return D.zeros(count, torch.float32)
        if a >= 0:
            return self._sources[0]._fetch(a, b)
        out = D.zeros(count, torch.float32)
        out[-a:] = self._sources[0]._fetch(0, b)
        return out

    def _values64(self, n0, n1):
        a, b = n0 - self._shift, n1 - self._shift               # source indices
        if b <= 0:
            return D.zeros(n1 - n0, torch.float64)
        if a >= 0:
            return self._sources[0]._fetch64(a, b)
        src = self._sources[0]._fetch64(0, b)
        return torch.cat((D.zeros(-a, torch.float64), src))


class EchoFilter(Filter):
    """Mixes `amount` echoes into the source, starting `after` seconds in; echo k is delayed
    k*delay seconds and scaled by decay**k (synthplayer.oscillators.EchoFilter).  A finite sum
    of delayed copies, so it is a gather over the rendered source, not a recurrence."""

    def __init__(self, source, after, amount, delay, decay):
        super().__init__([source])
        if decay < 1:
            amount = int(min(amount, log(0.000001, decay)))     # echoes nobody can hear
        if amount > C.ECHO_MAX:
            raise ValueError("at most %d echoes are supported" % C.ECHO_MAX)
        self._after = after
        self._amount = amount
        self._delay = delay
        self._decay = decay
        self.echo_duration = self._after + self._amount * self._delay
        self._n_after = int(self._samplerate * after)
        delays, gains = [], []
        amp, echo_delay = decay, delay
        for _ in range(amount):
            delays.append(int(self._samplerate * echo_delay))
            gains.append(amp)
            echo_delay += delay
            amp *= decay
        self._delays = (ctypes.c_int64 * max(1, amount))(*delays)
        self._gains = (ctypes.c_float * max(1, amount))(*gains)
        self._maxdelay = max(delays) if delays else 0

    def _render_range(self, n0, count):
        lo = max(0, min(n0, max(self._n_after, n0 - self._maxdelay)))
        src = self._sources[0]._fetch(lo, n0 + count)
        out = D.empty(count, torch.float32)
        C.call("sb_echo_f32", D.ptr(src), lo, self._n_after, self._amount, self._delays, self._gains,
               n0, count, D.ptr(out), D.stream())
        return out


class ClipFilter(Filter):
    """Clips the source to [minimum, maximum] (synthplayer.oscillators.ClipFilter)."""

    def __init__(self, source, minimum=-sys.float_info.max, maximum=sys.float_info.max):
        super().__init__([source])
        self.min = minimum
        self.max = maximum

    def _render_range(self, n0, count):
        src = self._sources[0]._fetch(n0, n0 + count)
        fmax = 3.4028234663852886e38
        out = D.empty(count, torch.float32)
        C.call("sb_clip_f32", D.ptr(src), max(-fmax, min(fmax, self.min)), max(-fmax, min(fmax, self.max)),
               count, D.ptr(out), D.stream())
        return out

    def _values64(self, a, b):
        return self._sources[0]._fetch64(a, b).clamp(self.min, self.max)


class AbsFilter(Filter):
    """Absolute value of the source (synthplayer.oscillators.AbsFilter)."""

    def __init__(self, source):
        super().__init__([source])

    def _render_range(self, n0, count):
        src = self._sources[0]._fetch(n0, n0 + count)
        out = D.empty(count, torch.float32)
        C.call("sb_abs_f32", D.ptr(src), count, D.ptr(out), D.stream())
        return out

    def _values64(self, a, b):
        return self._sources[0]._fetch64(a, b).abs()


class NullFilter(Filter):
    """Passes the source through unchanged (synthplayer.oscillators.NullFilter)."""

    def __init__(self, source):
        super().__init__([source])

    def _render_range(self, n0, count):
        return self._sources[0]._fetch(n0, n0 + count)

    def _values64(self, a, b):
        return self._sources[0]._fetch64(a, b)
