"""
ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product.

CPU restatement of synthplayer's ``Sample`` (module ``synthplayer.sample``): integer PCM
held as little-endian bytes, with mixing, resampling and amplitude operations.

PARITY: the *call sites* (which audioop routine each Sample method invokes, frame-index
rounding, fade formulas) are recalled from the published package — the source is absent
from /root/reference (README.md:1-2 only), so they are UNPINNED.  The *arithmetic* is
pinned: it is CPython 3.12.3's stdlib ``audioop``, executed directly when importable
(it is, in this image and on the GPU box), else oracle/audioop_port.py, which
tests/test_oracle_audioop.py checks against the real module.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""

import array
import itertools
import sys
import warnings

try:
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", DeprecationWarning)
        import audioop                      # the real thing: CPython Modules/audioop.c
    AUDIOOP_IS_STDLIB = True
except ImportError:                          # Python >= 3.13
    from . import audioop_port as audioop    # type: ignore
    AUDIOOP_IS_STDLIB = False

from . import synth_oracle as so


class Sample:
    """synthplayer.sample.Sample (recalled).  Most operations modify in place and return self."""

    norm_samplerate = so.NORM_SAMPLERATE
    norm_nchannels = so.NORM_NCHANNELS
    norm_samplewidth = so.NORM_SAMPLEWIDTH

    def __init__(self, name="", samplerate=0, nchannels=0, samplewidth=0):
        self.name = name
        self._locked = False
        self.samplerate = samplerate or self.norm_samplerate
        self.nchannels = nchannels or self.norm_nchannels
        self.samplewidth = samplewidth or self.norm_samplewidth
        self._frames = b""

    # -- construction ---------------------------------------------------------------
    @classmethod
    def from_raw_frames(cls, frames, samplewidth, samplerate, numchannels, name=""):
        assert 1 <= numchannels <= 2
        assert 2 <= samplewidth <= 4
        assert samplerate > 1
        s = cls(name=name, samplerate=samplerate, nchannels=numchannels, samplewidth=samplewidth)
        s._frames = bytes(frames)
        return s

    @classmethod
    def from_array(cls, array_or_list, samplerate, numchannels, name=""):
        assert 1 <= numchannels <= 2
        assert samplerate > 1
        if isinstance(array_or_list, list):
            try:
                array_or_list = cls.get_array(2, array_or_list)
            except OverflowError:
                array_or_list = cls.get_array(4, array_or_list)
        elif array_or_list.typecode not in "hil":
            raise ValueError("the sample values must be integer")
        frames = array_or_list.tobytes()
        if sys.byteorder == "big":
            frames = audioop.byteswap(frames, array_or_list.itemsize)
        return cls.from_raw_frames(frames, array_or_list.itemsize, samplerate, numchannels, name)

    @classmethod
    def from_osc_block(cls, block, samplerate, amplitude_scale=None, samplewidth=2):
        scale = amplitude_scale or (2 ** (8 * samplewidth - 1) - 1)
        ints = [int(v * scale) for v in block] if scale != 1.0 else [int(v) for v in block]
        return cls.from_array(cls.get_array(samplewidth, ints), samplerate, 1)

    @classmethod
    def from_oscillator(cls, osc, duration, samplerate=None, amplitude_scale=None, samplewidth=2):
        samplerate = samplerate or osc._samplerate
        values = so.render_floats(osc, int(samplerate * duration))
        return cls.from_osc_block(values, samplerate, amplitude_scale, samplewidth)

    @staticmethod
    def get_array(samplewidth, initializer=None):
        if samplewidth == 2:
            return array.array("h", initializer or [])
        if samplewidth == 4:
            return array.array("i" if array.array("i").itemsize == 4 else "l", initializer or [])
        raise ValueError("only samplewidth 2 or 4 are supported")

    # -- properties -----------------------------------------------------------------
    @property
    def duration(self):
        return len(self._frames) / self.samplerate / self.samplewidth / self.nchannels

    @property
    def maximum(self):
        return audioop.max(self._frames, self.samplewidth)

    def __len__(self):
        return len(self._frames) // self.samplewidth // self.nchannels

    def frame_idx(self, seconds):
        return self.nchannels * self.samplewidth * int(self.samplerate * seconds)

    def get_frame_array(self):
        return self.get_array(self.samplewidth, None if not self._frames else
                              array.array(self.get_array(self.samplewidth).typecode, self._frames))

    def view_frame_data(self):
        return memoryview(self._frames)

    def copy(self):
        c = Sample(self.name, self.samplerate, self.nchannels, self.samplewidth)
        c._frames = self._frames
        return c

    def lock(self):
        self._locked = True
        return self

    # -- rate / width / channels ----------------------------------------------------
    def resample(self, samplerate):
        assert not self._locked
        if samplerate == self.samplerate:
            return self
        self._frames = audioop.ratecv(self._frames, self.samplewidth, self.nchannels,
                                      self.samplerate, samplerate, None)[0]
        self.samplerate = samplerate
        return self

    def speed(self, speed):
        assert not self._locked
        assert speed > 0
        if speed == 1.0:
            return self
        rate = self.samplerate
        self._frames = audioop.ratecv(self._frames, self.samplewidth, self.nchannels,
                                      int(self.samplerate * speed), rate, None)[0]
        self.samplerate = rate
        return self

    def make_32bit(self, scale_amplitude=True):
        assert not self._locked
        self._frames = self.get_32bit_frames(scale_amplitude)
        self.samplewidth = 4
        return self

    def get_32bit_frames(self, scale_amplitude=True):
        if self.samplewidth == 4:
            return self._frames
        frames = audioop.lin2lin(self._frames, self.samplewidth, 4)
        if not scale_amplitude:
            frames = audioop.mul(frames, 4, 2 ** (8 * self.samplewidth) / 2 ** 32)
        return frames

    def make_16bit(self, maximize_amplitude=True):
        assert not self._locked
        assert self.samplewidth >= 2
        if maximize_amplitude:
            self.amplify_max()
        if self.samplewidth > 2:
            self._frames = audioop.lin2lin(self._frames, self.samplewidth, 2)
            self.samplewidth = 2
        return self

    def mono(self, left_factor=1.0, right_factor=1.0):
        assert not self._locked
        if self.nchannels == 1:
            return self
        if self.nchannels == 2:
            self._frames = audioop.tomono(self._frames, self.samplewidth, left_factor, right_factor)
            self.nchannels = 1
            return self
        raise ValueError("sample must be stereo or mono already")

    def left(self):
        assert self.nchannels == 2
        return self.mono(1.0, 0)

    def right(self):
        assert self.nchannels == 2
        return self.mono(0, 1.0)

    def stereo(self, left_factor=1.0, right_factor=1.0):
        assert not self._locked
        if self.nchannels == 2:
            # already stereo: split, scale each side, recombine (recalled)
            right = self.copy().right()._frames
            left = self.copy().left()._frames
            left = audioop.tostereo(left, self.samplewidth, left_factor, 0)
            right = audioop.tostereo(right, self.samplewidth, 0, right_factor)
            self._frames = audioop.add(left, right, self.samplewidth)
            return self
        if self.nchannels == 1:
            self._frames = audioop.tostereo(self._frames, self.samplewidth, left_factor, right_factor)
            self.nchannels = 2
            return self
        raise ValueError("sample must be mono or stereo already")

    def normalize(self):
        """synthplayer.sample.Sample.normalize (recalled): to the default rate, width, channels."""
        assert not self._locked
        self.resample(self.norm_samplerate)
        if self.samplewidth != self.norm_samplewidth:
            self._frames = audioop.lin2lin(self._frames, self.samplewidth, self.norm_samplewidth)
            self.samplewidth = self.norm_samplewidth
        if self.nchannels == 1 and self.norm_nchannels == 2:
            self._frames = audioop.tostereo(self._frames, self.samplewidth, 1, 1)
            self.nchannels = 2
        elif self.nchannels == 2 and self.norm_nchannels == 1:
            self.mono()
        return self

    def stereo_mix(self, other, other_channel, other_mix_factor=1.0, mix_at=0.0, other_seconds=None):
        """synthplayer.sample.Sample.stereo_mix (recalled)."""
        assert not self._locked
        assert other.nchannels == 1
        assert other.samplerate == self.samplerate
        assert other.samplewidth == self.samplewidth
        assert other_channel in ("L", "R")
        if self.nchannels == 1:
            if other_channel == "L":
                self.stereo(left_factor=0, right_factor=1)
            else:
                self.stereo(left_factor=1, right_factor=0)
        other = other.copy()
        if other_channel == "L":
            other.stereo(left_factor=other_mix_factor, right_factor=0)
        else:
            other.stereo(left_factor=0, right_factor=other_mix_factor)
        return self.mix_at(mix_at, other, other_seconds)

    def pan(self, panning=0.0, lfo=None):
        """(recalled, LOW confidence in the scaling: upstream may use (1-p)/2 and (1+p)/2.)"""
        assert not self._locked
        if lfo is not None:
            values = iter(lfo.generator())
            frames = self.get_frame_array()
            left = frames[0::2] if self.nchannels == 2 else frames
            right = frames[1::2] if self.nchannels == 2 else frames
            out = self.get_array(self.samplewidth, [0] * (2 * len(left)))
            for i in range(len(left)):
                p = next(values)
                out[2 * i] = int(left[i] * (1.0 - p))
                out[2 * i + 1] = int(right[i] * (1.0 + p))
            self._frames = out.tobytes()
            self.nchannels = 2
            return self
        assert -1 <= panning <= 1
        left_volume = (1.0 - panning) / 2
        right_volume = 1.0 - left_volume
        return self.stereo(left_volume * 2, right_volume * 2)

    # -- amplitude -------------------------------------------------------------------
    def amplify(self, factor):
        assert not self._locked
        self._frames = audioop.mul(self._frames, self.samplewidth, factor)
        return self

    def amplify_max(self):
        assert not self._locked
        max_amp = audioop.max(self._frames, self.samplewidth)
        max_target = 2 ** (8 * self.samplewidth - 1) - 2
        if max_amp > 0:
            factor = max_target / max_amp
            self._frames = audioop.mul(self._frames, self.samplewidth, factor)
        return self

    def at_volume(self, volume):
        c = self.copy()
        c.amplify(volume)
        return c

    def invert(self):
        assert not self._locked
        return self.amplify(-1)

    def bias(self, bias):
        assert not self._locked
        self._frames = audioop.bias(self._frames, self.samplewidth, bias)
        return self

    def reverse(self):
        assert not self._locked
        self._frames = audioop.reverse(self._frames, self.samplewidth)
        return self

    def fadeout(self, seconds, target_volume=0.0):
        assert not self._locked
        if not self._frames:
            return self
        seconds = min(seconds, self.duration)
        i = self.frame_idx(self.duration - seconds)
        begin = self._frames[:i]
        end = self._frames[i:]
        numsamples = len(end) / self.samplewidth
        decrease = 1.0 - target_volume
        sw = self.samplewidth
        getsample = audioop.getsample
        faded = self.get_array(sw, [int(getsample(end, sw, k) * (1.0 - k * decrease / numsamples))
                                    for k in range(int(numsamples))])
        end = faded.tobytes()
        if sys.byteorder == "big":
            end = audioop.byteswap(end, sw)
        self._frames = begin + end
        return self

    def fadein(self, seconds, start_volume=0.0):
        assert not self._locked
        if not self._frames:
            return self
        seconds = min(seconds, self.duration)
        i = self.frame_idx(seconds)
        begin = self._frames[:i]
        end = self._frames[i:]
        numsamples = len(begin) / self.samplewidth
        increase = 1.0 - start_volume
        sw = self.samplewidth
        getsample = audioop.getsample
        faded = self.get_array(sw, [int(getsample(begin, sw, k) * (k * increase / numsamples + start_volume))
                                    for k in range(int(numsamples))])
        begin = faded.tobytes()
        if sys.byteorder == "big":
            begin = audioop.byteswap(begin, sw)
        self._frames = begin + end
        return self

    def modulate_amp(self, modulation_source):
        assert not self._locked
        frames = self.get_frame_array()
        if isinstance(modulation_source, (Sample, list, array.array)):
            if isinstance(modulation_source, Sample):
                modulation_source = modulation_source.get_frame_array()
            biggest = max(max(modulation_source), abs(min(modulation_source)))
            actual = (v / biggest for v in itertools.cycle(modulation_source))
        elif isinstance(modulation_source, so.Oscillator):
            actual = iter(modulation_source.generator())
        else:
            actual = iter(modulation_source)
        for k in range(len(frames)):
            frames[k] = int(frames[k] * next(actual))
        self._frames = frames.tobytes()
        if sys.byteorder == "big":
            self._frames = audioop.byteswap(self._frames, self.samplewidth)
        return self

    # -- structure ---------------------------------------------------------------------
    def clip(self, start_seconds, end_seconds):
        assert not self._locked
        assert end_seconds >= start_seconds
        self._frames = self._frames[self.frame_idx(start_seconds):self.frame_idx(end_seconds)]
        return self

    def split(self, seconds):
        assert not self._locked
        end = self.frame_idx(seconds)
        if end != len(self._frames):
            chopped = self.copy()
            chopped._frames = self._frames[end:]
            self._frames = self._frames[:end]
            return chopped
        return Sample.from_raw_frames(b"", self.samplewidth, self.samplerate, self.nchannels)

    def add_silence(self, seconds, at_start=False):
        assert not self._locked
        required_extra = self.frame_idx(seconds)
        if at_start:
            self._frames = b"\0" * required_extra + self._frames
        else:
            self._frames += b"\0" * required_extra
        return self

    def join(self, other):
        assert not self._locked
        assert self.samplewidth == other.samplewidth
        assert self.samplerate == other.samplerate
        assert self.nchannels == other.nchannels
        self._frames += other._frames
        return self

    def delay(self, seconds, keep_length=False):
        assert not self._locked
        if seconds > 0:
            if keep_length:
                num_frames = len(self._frames)
                self.add_silence(seconds, at_start=True)
                self._frames = self._frames[:num_frames]
                return self
            return self.add_silence(seconds, at_start=True)
        if seconds < 0:
            seconds = -seconds
            if keep_length:
                num_frames = len(self._frames)
                self.add_silence(seconds)
                self._frames = self._frames[len(self._frames) - num_frames:]
                return self
            self._frames = self._frames[self.frame_idx(seconds):]
        return self

    # -- mixing -----------------------------------------------------------------------
    def mix(self, other, other_seconds=None, pad_shortest=True):
        assert not self._locked
        assert self.samplewidth == other.samplewidth
        assert self.samplerate == other.samplerate
        assert self.nchannels == other.nchannels
        frames1 = self._frames
        if other_seconds:
            frames2 = other._frames[:other.frame_idx(other_seconds)]
        else:
            frames2 = other._frames
        if pad_shortest:
            if len(frames1) < len(frames2):
                frames1 += b"\0" * (len(frames2) - len(frames1))
            elif len(frames2) < len(frames1):
                frames2 += b"\0" * (len(frames1) - len(frames2))
        self._frames = audioop.add(frames1, frames2, self.samplewidth)
        return self

    def mix_at(self, seconds, other, other_seconds=None):
        if seconds == 0.0:
            return self.mix(other, other_seconds)
        assert not self._locked
        assert self.samplewidth == other.samplewidth
        assert self.samplerate == other.samplerate
        assert self.nchannels == other.nchannels
        start_frame_idx = self.frame_idx(seconds)
        if other_seconds:
            other_frames = other._frames[:other.frame_idx(other_seconds)]
        else:
            other_frames = other._frames
        required_length = start_frame_idx + len(other_frames)
        if required_length > len(self._frames):
            self._frames += b"\0" * (required_length - len(self._frames))
        pre = self._frames[:start_frame_idx]
        to_mix = self._frames[start_frame_idx:start_frame_idx + len(other_frames)]
        post = self._frames[start_frame_idx + len(other_frames):]
        mixed = audioop.add(to_mix, other_frames, self.samplewidth)
        self._frames = pre + mixed + post
        return self

    def echo(self, length, amount, delay, decay):
        assert not self._locked
        if amount > 0:
            length = max(0, self.duration - length)
            echo = self.copy()
            echo._frames = self._frames[self.frame_idx(length):]
            echo_amp = decay
            for _ in range(amount):
                if echo_amp < 1.0 / (2 ** (8 * self.samplewidth - 1)):
                    break           # echoes nobody can hear
                length += delay
                echo = echo.copy().amplify(echo_amp)
                self.mix_at(length, echo)
                echo_amp *= decay
        return self

    def envelope(self, attack, decay, sustainlevel, release):
        assert not self._locked
        assert attack >= 0 and decay >= 0 and release >= 0
        assert 0 <= sustainlevel <= 1
        D = self.split(attack)      # self keeps the attack part
        S = D.split(decay)
        if sustainlevel < 1:
            S.amplify(sustainlevel)  # sustain level applied now, so that R gets it too
        R = S.split(S.duration - release)
        if attack > 0:
            self.fadein(attack)
        if decay > 0:
            D.fadeout(decay, sustainlevel)
        if release > 0:
            R.fadeout(release)
        self.join(D).join(S).join(R)
        return self


class WaveSynth:
    """synthplayer.synth.WaveSynth (recalled): the waveform methods build the oscillator with the
    integer scale folded into amplitude and bias, take int(samplerate*duration) values and store
    int(v) of each (truncation toward zero) in a mono Sample."""

    def __init__(self, samplerate=0, samplewidth=0):
        self.samplerate = samplerate or so.NORM_SAMPLERATE
        self.samplewidth = samplewidth or so.NORM_SAMPLEWIDTH
        if self.samplewidth not in (2, 4):
            raise ValueError("only samplewidth 2 or 4 supported")

    def _scale(self, freq, amplitude, bias):
        assert freq <= self.samplerate / 2
        assert 0 <= amplitude <= 1.0
        assert -1 <= bias <= 1.0
        return 2 ** (self.samplewidth * 8 - 1) - 1

    def _sample(self, osc, duration):
        values = so.render_floats(osc, int(self.samplerate * duration))
        return Sample.from_array(Sample.get_array(self.samplewidth, [int(v) for v in values]), self.samplerate, 1)

    def sine(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Sine(frequency, amplitude * k, phase, bias * k, fm_lfo=fm_lfo, samplerate=self.samplerate), duration)

    def square(self, frequency, duration, amplitude=0.75, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Square(frequency, amplitude * k, phase, bias * k, fm_lfo=fm_lfo, samplerate=self.samplerate), duration)

    def triangle(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Triangle(frequency, amplitude * k, phase, bias * k, fm_lfo=fm_lfo, samplerate=self.samplerate), duration)

    def sawtooth(self, frequency, duration, amplitude=0.75, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Sawtooth(frequency, amplitude * k, phase, bias * k, fm_lfo=fm_lfo, samplerate=self.samplerate), duration)

    def pulse(self, frequency, duration, amplitude=0.75, phase=0.0, bias=0.0, pulsewidth=0.1, fm_lfo=None, pwm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Pulse(frequency, amplitude * k, phase, bias * k, pulsewidth, fm_lfo=fm_lfo,
                                     pwm_lfo=pwm_lfo, samplerate=self.samplerate), duration)

    def harmonics(self, frequency, duration, harmonics, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Harmonics(frequency, harmonics, amplitude * k, phase, bias * k, fm_lfo=fm_lfo,
                                         samplerate=self.samplerate), duration)

    def square_h(self, frequency, duration, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.SquareH(frequency, num_harmonics, amplitude * k, phase, bias * k, fm_lfo=fm_lfo,
                                       samplerate=self.samplerate), duration)

    def sawtooth_h(self, frequency, duration, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.SawtoothH(frequency, num_harmonics, amplitude * k, phase, bias * k, fm_lfo=fm_lfo,
                                         samplerate=self.samplerate), duration)

    def semicircle(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Semicircle(frequency, amplitude * k, phase, bias * k, fm_lfo=fm_lfo, samplerate=self.samplerate), duration)

    def pointy(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        k = self._scale(frequency, amplitude, bias)
        return self._sample(so.Pointy(frequency, amplitude * k, phase, bias * k, fm_lfo=fm_lfo, samplerate=self.samplerate), duration)
