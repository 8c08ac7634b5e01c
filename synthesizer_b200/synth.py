"""
``WaveSynth`` with the API of ``synthplayer.synth.WaveSynth``: waveform methods that return a
``Sample`` (rendered on the GPU end to end: oscillator kernel -> float->PCM kernel) and
``*_gen`` variants that return the oscillator.  Upstream folds the integer scale into the
oscillator amplitude (amplitude * (2**(8*width-1)-1)) and truncates with int(); so does this.
(synthplayer/synth.py is absent from /root/reference; recalled from the published package.)
"""
from . import params
from .oscillators import (Harmonics, Linear, Oscillator, Pointy, Pulse, Sawtooth, SawtoothH, Semicircle, Sine, Square,
                          SquareH, Triangle, WhiteNoise)
from .sample import Sample

__all__ = ["WaveSynth", "key_num", "key_freq", "octave_notes", "note_freq"]

octave_notes = ["C", "C#", "D", "D#", "E", "F", "F#", "G", "G#", "A", "A#", "B"]


def key_num(note, octave):
    return (octave - 1) * 12 + octave_notes.index(note.upper()) + 4


def key_freq(key_number, a4=440.0):
    """Frequency of a note on a standard 88 key piano keyboard."""
    return 2 ** ((key_number - 49) / 12) * a4


def note_freq(note, octave=None, a4=440.0):
    if not octave:
        octave = int(note[-1:])
        note = note[:-1]
    return key_freq(key_num(note, octave), a4)


class WaveSynth:
    """Waveform sample synthesizer: sine, square, triangle, sawtooth, pulse, harmonics, square_h,
    sawtooth_h, semicircle, pointy, white noise, linear; optional FM LFO (the methods and their
    defaults as recalled of synthplayer.synth.WaveSynth).  Results are 16 or 32 bit integer Samples."""

    def __init__(self, samplerate=0, samplewidth=0):
        samplerate = samplerate or params.norm_samplerate
        samplewidth = samplewidth or params.norm_samplewidth
        if samplewidth not in (2, 4):
            raise ValueError("only samplewidth 2 or 4 supported")
        self.samplerate = samplerate
        self.samplewidth = samplewidth

    def to_sample(self, sample_array, fadeout=True):
        s = Sample.from_array(sample_array, self.samplerate, 1)
        return s.fadeout(0.1 if fadeout else 0)

    def _scale(self, freq, amplitude, bias):
        assert freq <= self.samplerate / 2        # Nyquist
        assert 0 <= amplitude <= 1.0
        assert -1 <= bias <= 1.0
        return 2 ** (self.samplewidth * 8 - 1) - 1

    def _render_sample(self, duration, osc):
        # the scale already sits in the oscillator's amplitude: int(v) only
        values = osc.render(int(self.samplerate * duration))
        return Sample._from_float(values, self.samplerate, 1.0, self.samplewidth)

    # ---- oscillator factories (the *_gen methods) ------------------------------------
    def sine_gen(self, frequency, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Sine(frequency, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo, samplerate=self.samplerate)

    def square_gen(self, frequency, amplitude=0.75, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Square(frequency, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo, samplerate=self.samplerate)

    def triangle_gen(self, frequency, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Triangle(frequency, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo, samplerate=self.samplerate)

    def sawtooth_gen(self, frequency, amplitude=0.75, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Sawtooth(frequency, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo, samplerate=self.samplerate)

    def pulse_gen(self, frequency, amplitude=0.75, phase=0.0, bias=0.0, pulsewidth=0.1, fm_lfo=None, pwm_lfo=None):
        assert 0 <= pulsewidth <= 1
        scale = self._scale(frequency, amplitude, bias)
        return Pulse(frequency, amplitude * scale, phase, bias * scale, pulsewidth, fm_lfo=fm_lfo,
                     pwm_lfo=pwm_lfo, samplerate=self.samplerate)

    def white_noise_gen(self, frequency, amplitude=0.9999, bias=0.0, noise_source=None):
        scale = self._scale(frequency, amplitude, bias)
        return WhiteNoise(frequency, amplitude * scale, bias * scale, samplerate=self.samplerate,
                          noise_source=noise_source)

    def harmonics_gen(self, frequency, harmonics, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Harmonics(frequency, harmonics, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo,
                         samplerate=self.samplerate)

    def square_h_gen(self, frequency, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return SquareH(frequency, num_harmonics, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo,
                       samplerate=self.samplerate)

    def sawtooth_h_gen(self, frequency, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return SawtoothH(frequency, num_harmonics, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo,
                         samplerate=self.samplerate)

    def semicircle_gen(self, frequency, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Semicircle(frequency, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo, samplerate=self.samplerate)

    def pointy_gen(self, frequency, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        scale = self._scale(frequency, amplitude, bias)
        return Pointy(frequency, amplitude * scale, phase, bias * scale, fm_lfo=fm_lfo, samplerate=self.samplerate)

    def linear_gen(self, startamp, increment=0.0, minamp=-1.0, maxamp=1.0):
        scale = 2 ** (self.samplewidth * 8 - 1) - 1
        return Linear(startamp * scale, increment * scale, minamp * scale, maxamp * scale,
                      samplerate=self.samplerate)

    # ---- Sample-returning methods ----------------------------------------------------
    def sine(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.sine_gen(frequency, amplitude, phase, bias, fm_lfo))

    def square(self, frequency, duration, amplitude=0.75, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.square_gen(frequency, amplitude, phase, bias, fm_lfo))

    def triangle(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.triangle_gen(frequency, amplitude, phase, bias, fm_lfo))

    def sawtooth(self, frequency, duration, amplitude=0.75, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.sawtooth_gen(frequency, amplitude, phase, bias, fm_lfo))

    def pulse(self, frequency, duration, amplitude=0.75, phase=0.0, bias=0.0, pulsewidth=0.1, fm_lfo=None,
              pwm_lfo=None):
        return self._render_sample(duration, self.pulse_gen(frequency, amplitude, phase, bias, pulsewidth,
                                                            fm_lfo, pwm_lfo))

    def white_noise(self, frequency, duration, amplitude=0.9999, bias=0.0, noise_source=None):
        return self._render_sample(duration, self.white_noise_gen(frequency, amplitude, bias, noise_source))

    def harmonics(self, frequency, duration, harmonics, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.harmonics_gen(frequency, harmonics, amplitude, phase, bias, fm_lfo))

    def square_h(self, frequency, duration, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.square_h_gen(frequency, num_harmonics, amplitude, phase, bias, fm_lfo))

    def sawtooth_h(self, frequency, duration, num_harmonics=16, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.sawtooth_h_gen(frequency, num_harmonics, amplitude, phase, bias, fm_lfo))

    def semicircle(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.semicircle_gen(frequency, amplitude, phase, bias, fm_lfo))

    def pointy(self, frequency, duration, amplitude=0.9999, phase=0.0, bias=0.0, fm_lfo=None):
        return self._render_sample(duration, self.pointy_gen(frequency, amplitude, phase, bias, fm_lfo))

    def linear(self, duration, startamp, increment=0.0, minamp=-1.0, maxamp=1.0):
        return self._render_sample(duration, self.linear_gen(startamp, increment, minamp, maxamp))
