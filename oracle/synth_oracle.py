"""
ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product.

CPU restatement (pure Python, serial, one sample at a time) of the render path of
irmen/synthesizer's ``synthplayer`` package: oscillators with FM, the filter chain,
and float -> integer PCM conversion.

PARITY UNPINNED.  /root/reference holds only a README that redirects to Codeberg
(/root/reference/README.md:1-2); ``synthplayer/oscillators.py``, ``sample.py`` and
``synth.py`` are absent offline and there is no network.  No reference file:line can be
cited, no golden vector from the reference's own tests exists, and the reference cannot
be executed here.  Everything below restates the *published* algorithm of the public
``synthplayer`` package (PyPI ``synthplayer`` 2.x, module ``synthplayer.oscillators``)
as the author of this file recalls it; each function names the upstream class/method it
follows.  When the source is mounted, every function here must be re-checked and this
header changed.  The one piece that IS pinned is the integer PCM arithmetic of ``Sample``
(see sample_oracle.py): upstream delegates it to CPython's stdlib ``audioop``, which is
importable in this image (CPython 3.12.3) and is executed directly.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product (synthesizer_b200/) never does.

Design notes that bear on parity (answers to SURVEY.md K1..K10, as recalled):
  K1  float->int: multiply by 2**(8*width-1)-1, then int() (truncate toward zero); values
      outside the integer range raise OverflowError in array('h') — no clipping.
  K2  block size params.norm_osc_blocksize = 512; block boundaries never change values.
  K3  every oscillator advances ``t += increment`` serially in fp64 (it drifts); FM enters
      as freq = f0*(1+lfo) with a running ``phase_correction += (freq_prev-freq)*t``.
  K6  EnvelopeFilter compares a serially accumulated ``time`` with the segment end times
      and accumulates ``amp += amp_change``; one extra sample at the end of release if
      amp is still > 0.
  K7  Delay/Echo: int(samplerate*seconds) samples; EchoFilter is a finite sum of delayed,
      decayed copies (FIR), not a feedback loop.
  K8  formulas at discontinuities are the fp64 expressions below, evaluated on the
      drifting t; which side of an edge a sample falls on follows from those roundings.
  K9  WhiteNoise draws random.uniform(-amp, amp) once per held value; tests inject the
      uniform draws instead (``noise_source``).
"""

import itertools
import random
import sys
from math import pi, sin, cos, copysign, fabs, floor, log, sqrt

NORM_SAMPLERATE = 44100
NORM_SAMPLEWIDTH = 2
NORM_NCHANNELS = 2
NORM_OSC_BLOCKSIZE = 512


class Oscillator:
    """synthplayer.oscillators.Oscillator (recalled): base class; blocks() yields lists of
    NORM_OSC_BLOCKSIZE floats forever (or until a filter stops), generator() single floats."""

    def __init__(self, source=None, samplerate=None):
        self._samplerate = samplerate or source._samplerate
        self._source = source

    def generator(self):
        raise NotImplementedError

    def blocks(self):
        gen = self.generator()
        while True:
            block = list(itertools.islice(gen, NORM_OSC_BLOCKSIZE))
            if not block:
                return
            yield block

    def __iter__(self):
        return self.blocks()


def _fm_iter(fm_lfo):
    return iter(fm_lfo.generator()) if fm_lfo is not None else itertools.repeat(0.0)


class Sine(Oscillator):
    """synthplayer.oscillators.Sine.blocks (recalled): phase in radians,
    increment = 2*pi/samplerate, value = sin(t*freq + phase_correction)*amplitude + bias."""

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self.frequency = frequency
        self.amplitude = amplitude
        self.bias = bias
        self._phase = phase
        self._fm_lfo = fm_lfo

    def _phases(self):
        fm = _fm_iter(self._fm_lfo)
        phase_correction = self._phase * 2 * pi
        freq_previous = self.frequency
        increment = 2.0 * pi / self._samplerate
        t = 0.0
        frequency = self.frequency
        while True:
            freq = frequency * (1.0 + next(fm))
            phase_correction += (freq_previous - freq) * t
            freq_previous = freq
            yield t * freq + phase_correction
            t += increment

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for x in self._phases():
            yield sin(x) * amplitude + bias


class _CycleOsc(Oscillator):
    """Common driver of the waveforms whose phase is counted in cycles
    (increment = 1/samplerate, tt = t*freq + phase_correction)."""

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self.frequency = frequency
        self.amplitude = amplitude
        self.bias = bias
        self._phase = phase
        self._fm_lfo = fm_lfo

    def _tt(self):
        fm = _fm_iter(self._fm_lfo)
        phase_correction = self._phase
        freq_previous = self.frequency
        increment = 1.0 / self._samplerate
        t = 0.0
        frequency = self.frequency
        while True:
            freq = frequency * (1.0 + next(fm))
            phase_correction += (freq_previous - freq) * t
            freq_previous = freq
            yield t * freq + phase_correction
            t += increment


class Triangle(_CycleOsc):
    """synthplayer.oscillators.Triangle.blocks (recalled)."""

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for tt in self._tt():
            yield 4.0 * amplitude * (fabs((tt + 0.75) % 1.0 - 0.5) - 0.25) + bias


class Square(_CycleOsc):
    """synthplayer.oscillators.Square.blocks (recalled): sign from int(tt*2) % 2."""

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for tt in self._tt():
            yield (-amplitude if int(tt * 2) % 2 else amplitude) + bias


class Sawtooth(_CycleOsc):
    """synthplayer.oscillators.Sawtooth.blocks (recalled)."""

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for tt in self._tt():
            yield bias + amplitude * 2.0 * (tt - floor(0.5 + tt))


class Pulse(_CycleOsc):
    """synthplayer.oscillators.Pulse.blocks (recalled): +amp while (tt % 1.0) < pulsewidth;
    the pulse width can be driven by a second LFO (pwm_lfo), clamped to (0, 1)."""

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, pulsewidth=0.1,
                 fm_lfo=None, pwm_lfo=None, samplerate=0):
        assert 0 <= pulsewidth <= 1
        super().__init__(frequency, amplitude, phase, bias, fm_lfo, samplerate)
        self.pulsewidth = pulsewidth
        self._pwm_lfo = pwm_lfo

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        epsilon = sys.float_info.epsilon
        pwm = iter(self._pwm_lfo.generator()) if self._pwm_lfo is not None \
            else itertools.repeat(self.pulsewidth)
        for tt in self._tt():
            pw = next(pwm)
            if pw <= 0.0:
                pw = epsilon
            elif pw >= 1.0:
                pw = 1.0 - epsilon
            yield (amplitude if tt % 1.0 < pw else -amplitude) + bias


class Semicircle(_CycleOsc):
    """synthplayer.oscillators.Semicircle (recalled, LOW confidence in the exact expression):
    ft = tt - floor(tt) - 0.5; value = sqrt(0.25 - ft*ft) * (2*amplitude) + bias."""

    def generator(self):
        amplitude2, bias = self.amplitude * 2.0, self.bias
        for tt in self._tt():
            ft = tt - floor(tt) - 0.5
            yield sqrt(0.25 - ft * ft) * amplitude2 + bias


class Pointy(Sine):
    """synthplayer.oscillators.Pointy (recalled, LOW confidence in the exact expression): phase in
    radians like Sine; vv = 1 - |cos x|; value = copysign(vv*vv, sin x) * amplitude + bias."""

    def generator(self):
        for x in self._phases():
            vv = 1.0 - fabs(cos(x))
            yield copysign(vv * vv, sin(x)) * self.amplitude + self.bias


class WhiteNoise(Oscillator):
    """synthplayer.oscillators.WhiteNoise.blocks (recalled): one random.uniform(-amp, amp)
    draw is held for int(samplerate/frequency) samples.  ``noise_source`` (an iterable of
    uniform draws in [0,1)) replaces the random module so that both sides of a parity test
    consume the same numbers: value = -amp + 2*amp*u, which is what random.uniform computes."""

    def __init__(self, frequency, amplitude=1.0, bias=0.0, samplerate=0, noise_source=None):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self.frequency = frequency
        self.amplitude = amplitude
        self.bias = bias
        self._noise_source = noise_source

    def generator(self):
        cycles = int(self._samplerate / self.frequency)
        if cycles < 1:
            raise ValueError("whitenoise frequency cannot be bigger than the sample rate")
        amplitude, bias = self.amplitude, self.bias
        src = iter(self._noise_source) if self._noise_source is not None else None
        while True:
            u = next(src) if src is not None else random.random()
            value = (-amplitude + (amplitude - (-amplitude)) * u) + bias
            for _ in range(cycles):
                yield value


class Linear(Oscillator):
    """synthplayer.oscillators.Linear (recalled): a ramp that is clamped to [min, max]."""

    def __init__(self, startlevel, increment=0.0, min_value=-1.0, max_value=1.0, samplerate=0):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self.value = startlevel
        self.increment = increment
        self.min_value = min_value
        self.max_value = max_value

    def generator(self):
        value = self.value
        while True:
            yield value
            value = min(self.max_value, max(self.min_value, value + self.increment))


class FastSine(Oscillator):
    """synthplayer.oscillators.FastSine (recalled): t starts at phase*2*pi and advances by
    2*pi*frequency/samplerate; no FM."""

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, samplerate=0):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self._frequency = frequency
        self._phase = phase
        self.amplitude = amplitude
        self.bias = bias

    def generator(self):
        rate = self._samplerate / self._frequency
        increment = 2.0 * pi / rate
        t = self._phase * 2.0 * pi
        amplitude, bias = self.amplitude, self.bias
        while True:
            yield sin(t) * amplitude + bias
            t += increment


class _FastCycle(Oscillator):
    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, samplerate=0):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self._frequency = frequency
        self._phase = phase
        self.amplitude = amplitude
        self.bias = bias

    def _t(self):
        freq = self._frequency
        t = self._phase / freq
        increment = 1.0 / self._samplerate
        while True:
            yield t * freq
            t += increment


class FastTriangle(_FastCycle):
    """synthplayer.oscillators.FastTriangle (recalled)."""

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for tt in self._t():
            yield 4.0 * amplitude * (fabs((tt + 0.75) % 1.0 - 0.5) - 0.25) + bias


class FastSquare(_FastCycle):
    """synthplayer.oscillators.FastSquare (recalled)."""

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for tt in self._t():
            yield (-amplitude if int(tt * 2) % 2 else amplitude) + bias


class FastSawtooth(_FastCycle):
    """synthplayer.oscillators.FastSawtooth (recalled)."""

    def generator(self):
        amplitude, bias = self.amplitude, self.bias
        for tt in self._t():
            yield bias + amplitude * 2.0 * (tt - floor(0.5 + tt))


class FastPulse(_FastCycle):
    """synthplayer.oscillators.FastPulse (recalled): fixed pulse width, no FM/PWM."""

    def __init__(self, frequency, amplitude=1.0, phase=0.0, bias=0.0, pulsewidth=0.1, samplerate=0):
        assert 0 <= pulsewidth <= 1
        super().__init__(frequency, amplitude, phase, bias, samplerate)
        self._pulsewidth = pulsewidth

    def generator(self):
        amplitude, bias, pw = self.amplitude, self.bias, self._pulsewidth
        for tt in self._t():
            yield (amplitude if tt % 1.0 < pw else -amplitude) + bias


class Harmonics(Oscillator):
    """synthplayer.oscillators.Harmonics (recalled): sum of sin(k*q)*a_k over the harmonics
    below Nyquist, q = t*freq + phase_correction in radians, FM like Sine."""

    def __init__(self, frequency, harmonics, amplitude=1.0, phase=0.0, bias=0.0, fm_lfo=None, samplerate=0):
        super().__init__(samplerate=samplerate or NORM_SAMPLERATE)
        self.frequency = frequency
        self.amplitude = amplitude
        self.bias = bias
        self._phase = phase
        self._fm_lfo = fm_lfo
        nyquist = self._samplerate / 2.0
        self.harmonics = [(n, a) for n, a in harmonics if n * frequency <= nyquist]

    def generator(self):
        fm = _fm_iter(self._fm_lfo)
        phase_correction = self._phase * 2.0 * pi
        freq_previous = self.frequency
        increment = 2.0 * pi / self._samplerate
        t = 0.0
        while True:
            h = 0.0
            freq = self.frequency * (1.0 + next(fm))
            phase_correction += (freq_previous - freq) * t
            freq_previous = freq
            q = t * freq + phase_correction
            for k, amp in self.harmonics:
                h += sin(q * k) * amp
            yield h * self.amplitude + self.bias
            t += increment


class SquareH(Harmonics):
    """synthplayer.oscillators.SquareH (recalled): odd harmonics with amplitude 1/n."""

    def __init__(self, frequency, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0,
                 fm_lfo=None, samplerate=0):
        harmonics = [(n, 1.0 / n) for n in range(1, num_harmonics * 2, 2)]
        super().__init__(frequency, harmonics, amplitude, phase, bias, fm_lfo, samplerate)


class SawtoothH(Harmonics):
    """synthplayer.oscillators.SawtoothH (recalled): all harmonics 1/n, inverted, phase+0.5."""

    def __init__(self, frequency, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0,
                 fm_lfo=None, samplerate=0):
        harmonics = [(n, 1.0 / n) for n in range(1, num_harmonics + 1)]
        super().__init__(frequency, harmonics, -amplitude, phase + 0.5, bias, fm_lfo, samplerate)


# ----------------------------------------------------------------------------- filters

class Filter(Oscillator):
    def __init__(self, sources):
        self._sources = list(sources)
        super().__init__(self._sources[0])


def _as_gen(src):
    return iter(src.generator()) if isinstance(src, Oscillator) else iter(src)


class EnvelopeFilter(Filter):
    """synthplayer.oscillators.EnvelopeFilter (recalled): ADSR in seconds, sustain_level a
    factor; after release the output is 0.0 forever unless stop_at_end."""

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

    def generator(self):
        oscillator = _as_gen(self._sources[0])
        time = 0.0
        end_time_decay = self._attack + self._decay
        end_time_sustain = end_time_decay + self._sustain
        end_time_release = end_time_sustain + self._release
        increment = 1.0 / self._samplerate
        try:
            if self._attack:
                amp_change = 1.0 / self._attack * increment
                amp = 0.0
                while time < self._attack:
                    yield next(oscillator) * amp
                    amp += amp_change
                    time += increment
            if self._decay:
                amp = 1.0
                amp_change = (self._sustain_level - 1.0) / self._decay * increment
                while time < end_time_decay:
                    yield next(oscillator) * amp
                    amp += amp_change
                    time += increment
            while time < end_time_sustain:
                yield next(oscillator) * self._sustain_level
                time += increment
            if self._release:
                amp = self._sustain_level
                amp_change = (-self._sustain_level) / self._release * increment
                while time < end_time_release:
                    yield next(oscillator) * amp
                    amp += amp_change
                    time += increment
                if amp > 0.0:
                    yield next(oscillator) * amp
            if not self._stop_at_end:
                while True:
                    yield 0.0
        except StopIteration:
            return


class MixingFilter(Filter):
    """synthplayer.oscillators.MixingFilter (recalled): per-sample sum of all sources."""

    def __init__(self, *sources):
        super().__init__(sources)

    def generator(self):
        gens = [_as_gen(s) for s in self._sources]
        try:
            while True:
                yield sum([next(g) for g in gens])
        except StopIteration:
            return


class AmpModulationFilter(Filter):
    """synthplayer.oscillators.AmpModulationFilter (recalled): source * modulator."""

    def __init__(self, source, modulator):
        super().__init__([source])
        self._modulator = modulator

    def generator(self):
        src = _as_gen(self._sources[0])
        mod = _as_gen(self._modulator)
        try:
            while True:
                yield next(src) * next(mod)
        except StopIteration:
            return


class DelayFilter(Filter):
    """synthplayer.oscillators.DelayFilter (recalled): int(samplerate*seconds) zeros in front,
    or that many samples skipped for a negative delay."""

    def __init__(self, source, seconds):
        super().__init__([source])
        self._seconds = seconds

    def generator(self):
        src = _as_gen(self._sources[0])
        if self._seconds < 0.0:
            amount = int(-self._samplerate * self._seconds)
            for _ in range(amount):
                next(src)
        else:
            for _ in range(int(self._samplerate * self._seconds)):
                yield 0.0
        yield from src


class EchoFilter(Filter):
    """synthplayer.oscillators.EchoFilter (recalled): after ``after`` seconds the stream is
    split; echo k (k = 1..amount) is the rest of the stream delayed by k*delay seconds and
    scaled by decay**k; all are summed with the dry stream."""

    def __init__(self, source, after, amount, delay, decay):
        super().__init__([source])
        if decay < 1:
            amount = int(min(amount, log(0.000001, decay)))   # echoes nobody can hear
        self._after = after
        self._amount = amount
        self._delay = delay
        self._decay = decay
        self.echo_duration = self._after + self._amount * self._delay

    def generator(self):
        oscillator = _as_gen(self._sources[0])
        for _ in range(int(self._samplerate * self._after)):
            yield next(oscillator)
        tees = itertools.tee(oscillator, self._amount + 1)
        streams = [tees[0]]
        amp = self._decay
        echo_delay = self._delay
        for tee in tees[1:]:
            streams.append(self._delayed_scaled(tee, int(self._samplerate * echo_delay), amp))
            echo_delay += self._delay
            amp *= self._decay
        try:
            while True:
                yield sum([next(s) for s in streams])
        except StopIteration:
            return

    @staticmethod
    def _delayed_scaled(stream, nzeros, amp):
        for _ in range(nzeros):
            yield 0.0
        for v in stream:
            yield v * amp


class ClipFilter(Filter):
    """synthplayer.oscillators.ClipFilter (recalled): max(min(v, maximum), minimum)."""

    def __init__(self, source, minimum=-sys.float_info.max, maximum=sys.float_info.max):
        super().__init__([source])
        self.min = minimum
        self.max = maximum

    def generator(self):
        vmin, vmax = self.min, self.max
        for v in _as_gen(self._sources[0]):
            yield max(min(v, vmax), vmin)


class AbsFilter(Filter):
    """synthplayer.oscillators.AbsFilter (recalled)."""

    def __init__(self, source):
        super().__init__([source])

    def generator(self):
        for v in _as_gen(self._sources[0]):
            yield fabs(v)


class NullFilter(Filter):
    """synthplayer.oscillators.NullFilter (recalled): passes the source through."""

    def __init__(self, source):
        super().__init__([source])

    def generator(self):
        yield from _as_gen(self._sources[0])


# ----------------------------------------------------------------------------- float -> PCM

def render_floats(osc, nframes):
    """First ``nframes`` values of an oscillator (shorter if a filter stops early)."""
    return list(itertools.islice(osc.generator(), nframes))


def to_pcm(values, samplewidth=2, scale=None):
    """WaveSynth / Sample.from_osc_block (recalled): v*scale then int() — truncation toward
    zero; out-of-range values are an OverflowError (array('h') would raise), never clipped."""
    if scale is None:
        scale = 2 ** (8 * samplewidth - 1) - 1
    lo, hi = -(2 ** (8 * samplewidth - 1)), 2 ** (8 * samplewidth - 1) - 1
    out = []
    for v in values:
        i = int(v * scale) if scale != 1.0 else int(v)
        if i < lo or i > hi:
            raise OverflowError("sample value out of range for samplewidth %d" % samplewidth)
        out.append(i)
    return out
