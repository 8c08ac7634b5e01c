"""
Device plumbing: torch supplies device memory and streams, nothing else.  Every computation
goes through the C ABI (``_cabi``).  Without a CUDA device these helpers raise: the package
has no CPU path.
"""
import ctypes
import math
from ctypes import c_void_p

import torch

from . import _cabi as C

_TB_MAX_N = 1 << 40          # the time base is built once per increment for 2^40 samples


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("synthesizer_b200 renders on a CUDA GPU (B200, sm_100a); "
                           "no GPU is visible and there is no CPU fallback")


def device():
    require_cuda()
    return torch.device("cuda", torch.cuda.current_device())


def stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def empty(n, dtype):
    return torch.empty(int(n), dtype=dtype, device=device())


def zeros(n, dtype):
    return torch.zeros(int(n), dtype=dtype, device=device())


class Timebase:
    """Host copy of the exact time base for one increment, plus lazily made device copies."""

    def __init__(self, increment, t0=0.0):
        self.increment = increment
        self.t0 = t0
        self.segs = (C.TimeSeg * C.TIMEBASE_MAX_SEGS)()
        self.nsegs = C.call("sb_timebase_build_from", t0, increment, _TB_MAX_N, self.segs, C.TIMEBASE_MAX_SEGS)
        self._dev = {}

    def at(self, n):
        return C.lib.sb_timebase_at(self.segs, self.nsegs, int(n))

    def first_ge(self, n_from, limit):
        return C.lib.sb_timebase_first_ge(self.segs, self.nsegs, int(n_from), float(limit))

    def dev(self):
        """Device pointer (as tensor) of the segment table on the current device."""
        idx = torch.cuda.current_device()
        t = self._dev.get(idx)
        if t is None:
            raw = ctypes.string_at(ctypes.addressof(self.segs), ctypes.sizeof(C.TimeSeg) * self.nsegs)
            t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(device())
            self._dev[idx] = t
        return t


_timebases = {}


def timebase(increment, t0=0.0):
    key = (increment, t0)
    tb = _timebases.get(key)
    if tb is None:
        if len(_timebases) > 4096:          # Fast* oscillators make one per (frequency, phase)
            _timebases.clear()
        tb = _timebases[key] = Timebase(increment, t0)
    return tb


def timebase_rad(samplerate):
    """increment = 2*pi/samplerate, the step of upstream's Sine / Harmonics."""
    return timebase(2.0 * math.pi / samplerate)


def timebase_cyc(samplerate):
    """increment = 1/samplerate, the step of the other waveforms and of EnvelopeFilter."""
    return timebase(1.0 / samplerate)


def build_envelope(samplerate, attack, decay, sustain, sustain_level, release):
    env = C.Envelope()
    C.call("sb_envelope_build", float(samplerate), float(attack), float(decay), float(sustain),
           float(sustain_level), float(release), ctypes.byref(env))
    return env


def launch_count():
    return int(C.lib.sb_launch_count())
