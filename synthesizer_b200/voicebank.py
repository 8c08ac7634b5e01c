"""
Polyphonic voice banks: many voices of the shape

    EnvelopeFilter( Carrier(fm_lfo=Modulator) ) * (left gain, right gain)

summed to one stereo stream — what upstream expresses as a MixingFilter over
AmpModulationFilter/EnvelopeFilter/oscillator chains, evaluated here by ONE fused kernel
(sb_bank_render_f32) plus a fold of the per-group partial mixes.  Voices shard across GPUs by
voice; the partial mixes are summed with one NCCL reduce and converted to PCM once, on rank 0.

Backend API (no upstream class of this name): upstream reaches the same result through its
oscillator classes; `VoiceBank.from_oscillators` accepts exactly those objects.
"""
import ctypes

import torch

from . import _cabi as C
from . import _device as D
from .oscillators import EnvelopeFilter, Harmonics, _WaveOscillator

__all__ = ["VoiceBank", "shard_range"]


def shard_range(nvoices, rank, world_size):
    """Contiguous, balanced slice [begin, end) of the voices that `rank` renders."""
    base, extra = divmod(nvoices, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def _voice_from(osc, gain_l, gain_r):
    v = C.Voice()
    env = None
    if isinstance(osc, EnvelopeFilter):
        if osc._stop_at_end:
            raise ValueError("voices with stop_at_end envelopes must use the general render path")
        env, osc = osc._env, osc._sources[0]
    if not isinstance(osc, _WaveOscillator) or isinstance(osc, Harmonics) or osc._pwm_lfo is not None \
            or osc._wave > C.WAVE_PULSE:
        raise ValueError("a bank voice is a Sine/Triangle/Square/Sawtooth/Pulse carrier without PWM")
    v.c_wave, v.c_freq, v.c_phase = osc._wave, osc.frequency, osc._phase
    v.c_amp, v.c_bias, v.c_pw = osc.amplitude, osc.bias, osc.pulsewidth
    lfo = osc._fm_lfo
    if lfo is None:
        v.m_wave = C.WAVE_NONE
    else:
        if not isinstance(lfo, _WaveOscillator) or isinstance(lfo, Harmonics) or lfo._fm_lfo is not None \
                or lfo._pwm_lfo is not None or lfo._wave > C.WAVE_PULSE:
            raise ValueError("the FM modulator of a bank voice must be a plain waveform oscillator")
        if lfo._samplerate != osc._samplerate:
            raise ValueError("carrier and modulator sample rates differ")
        v.m_wave, v.m_freq, v.m_phase = lfo._wave, lfo.frequency, lfo._phase
        v.m_amp, v.m_bias, v.m_pw = lfo.amplitude, lfo.bias, lfo.pulsewidth
    v.gain_l, v.gain_r = gain_l, gain_r
    if env is not None:
        v.env = env
    else:
        v.env.enabled = 0
    return v, osc._samplerate


class VoiceBank:
    """A set of voices rendered and mixed by the fused bank kernel."""

    def __init__(self, samplerate):
        self.samplerate = samplerate
        self._voices = []
        self._dev = None
        self._last = None
        self._scratch = None

    def __len__(self):
        return len(self._voices)

    def add(self, osc, gain_left=1.0, gain_right=1.0):
        v, sr = _voice_from(osc, gain_left, gain_right)
        if sr != self.samplerate:
            raise ValueError("voice sample rate %s differs from the bank's %s" % (sr, self.samplerate))
        self._voices.append(v)
        self._dev = None
        return self

    @classmethod
    def from_oscillators(cls, oscillators, gains=None, samplerate=None):
        oscillators = list(oscillators)
        bank = cls(samplerate or oscillators[0]._samplerate)
        for i, osc in enumerate(oscillators):
            gl, gr = gains[i] if gains is not None else (1.0, 1.0)
            bank.add(osc, gl, gr)
        return bank

    @classmethod
    def from_voice_array(cls, voices, samplerate):
        """From a ctypes array / list of _cabi.Voice (already built, e.g. by bench.py)."""
        bank = cls(samplerate)
        bank._voices = list(voices)
        return bank

    def invalidate(self):
        """Forget the uploaded copy of the voices (call after changing a voice in place)."""
        self._dev = None

    def voice_array(self, begin=0, end=None):
        sel = self._voices[begin:end]
        return (C.Voice * len(sel))(*sel)

    def _upload(self, begin, end):
        arr = self.voice_array(begin, end)
        raw = ctypes.string_at(ctypes.addressof(arr), ctypes.sizeof(arr))
        return torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(D.device()), len(arr)

    def count_classes(self, begin=0, end=None):
        """(generic, smooth, edged, scan) voice counts of [begin, end): how many voices each kernel
        of a render works on (sb_bank_count_classes; the hints of sb_bank_plan_hint)."""
        arr = self.voice_array(begin, end)
        if not len(arr):
            return (0, 0, 0, 0)
        counts = (ctypes.c_int * 4)()
        C.call("sb_bank_count_classes", ctypes.byref(arr), len(arr), float(self.samplerate), counts)
        return tuple(counts)

    def count_generic(self, begin=0, end=None):
        """Voices of [begin, end) that need the generic (per-sample fp64) kernel."""
        return self.count_classes(begin, end)[0]

    def count_scan(self, begin=0, end=None):
        """Voices of [begin, end) on the fast path with a Triangle / Square / Sawtooth / Pulse
        modulator (class 2: their FM sum comes from a tile pre-pass)."""
        return self.count_classes(begin, end)[3]

    def count_edged(self, begin=0, end=None):
        """Class-1 voices of [begin, end) with a Square / Sawtooth / Pulse carrier (the second fast kernel)."""
        return self.count_classes(begin, end)[2]

    def render(self, nframes, start=0, begin=0, end=None, voices_dev=None, generic_voices=None,
               edged_voices=None, scan_voices=None, smooth_voices=None, out=None):
        """Interleaved stereo float32 mix (2*nframes values) of voices [begin, end) for frames
        [start, start+nframes), as a CUDA tensor."""
        D.require_cuda()
        if voices_dev is None:
            # the uploaded voices and their class counts are kept until a voice is added
            # (chunked renders - blocks(), PcmStreamer - come back here for every chunk)
            if self._dev is None:
                self._dev = {}
            key = (begin, end)
            if key not in self._dev:
                self._dev[key] = self._upload(begin, end) + self.count_classes(begin, end)
            voices_dev, nv, generic_voices, smooth_voices, edged_voices, scan_voices = self._dev[key]
        else:
            nv = voices_dev.numel() // ctypes.sizeof(C.Voice)
        if out is not None and (out.dtype != torch.float32 or out.numel() != 2 * nframes or not out.is_cuda):
            raise ValueError("out must be a CUDA float32 tensor of 2*nframes values")
        mix = out if out is not None else D.empty(2 * nframes, torch.float32)
        if nv == 0:
            return mix.zero_()
        plan = C.BankPlan()
        sms = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
        tr, tc = D.timebase_rad(self.samplerate), D.timebase_cyc(self.samplerate)
        C.call("sb_bank_plan_make", nv, nframes, start, sms, tr.nsegs, ctypes.byref(plan))
        unknown = lambda x: -1 if x is None else x
        C.call("sb_bank_plan_hint", ctypes.byref(plan), unknown(generic_voices), unknown(smooth_voices),
               unknown(edged_voices), unknown(scan_voices))
        # the scratch is kept between renders (a bank is usually rendered again and again: chunks,
        # steps); it only grows
        need = (max(1, plan.carry_bytes // 8), max(1, plan.partial_bytes // 4), torch.cuda.current_device())
        if self._scratch is None or self._scratch[0].numel() < need[0] or self._scratch[1].numel() < need[1] \
                or self._scratch[2] != need[2]:
            self._scratch = None                         # release before growing
            self._scratch = (D.empty(need[0], torch.float64), D.empty(need[1], torch.float32), need[2])
        carry, partial = self._scratch[0], self._scratch[1]
        C.call("sb_bank_render_f32", D.ptr(voices_dev), ctypes.byref(plan), D.ptr(tr.dev()), tr.nsegs,
               D.ptr(tc.dev()), tc.nsegs, start, float(self.samplerate), D.ptr(carry), D.ptr(partial),
               D.ptr(mix), D.stream())
        self._last = (plan, carry)
        return mix

    def check_last_render(self):
        """Waits for the last render and returns how many of its voices were left to the generic
        kernels; raises if the hints it was given said none would be (sb_bank_render_check):
        those voices are then missing from the mix.  Only needed with hand-made hints."""
        plan, carry = self._last
        n = ctypes.c_int(0)
        C.call("sb_bank_render_check", ctypes.byref(plan), D.ptr(carry), ctypes.byref(n), D.stream())
        return n.value

    def render_sharded(self, nframes, start=0, group=None, out=None):
        """Voice-sharded render: this rank renders its slice of the voices; the partial mixes
        are summed on rank 0 with one reduce.  Returns the full mix on rank 0 (the partial one
        elsewhere).  Without an initialised process group this is a plain render.  `out`: a CUDA
        float32 tensor of 2*nframes values to render into (a caller that renders repeatedly keeps
        one, and the reduce then always works on the same buffer)."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return self.render(nframes, start, out=out)
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        b, e = shard_range(len(self._voices), rank, world)
        mix = self.render(nframes, start, b, e, out=out)
        dist.reduce(mix, dst=0, op=dist.ReduceOp.SUM, group=group)
        return mix

    def to_sample(self, duration, scale=None, sharded=False):
        """Render `duration` seconds to a stereo 16-bit Sample (int(v*scale), truncation)."""
        from .sample import Sample
        nframes = int(self.samplerate * duration)
        mix = self.render_sharded(nframes) if sharded else self.render(nframes)
        s = Sample._from_float(mix, self.samplerate, scale, 2)
        s.nchannels = 2
        return s
