"""
Chunked PCM output of an oscillator graph: the step after the render path (what upstream's
playback / streaming code consumes).  Backend API — upstream reaches the same bytes by pulling
``osc.blocks()`` and converting each block on the CPU.

``PcmStreamer`` renders ahead on the GPU: chunk k+1 is rendered and converted to PCM on a side
stream, and its device->host copy into a pinned ring buffer is in flight, while the caller is
still consuming chunk k.  Nothing here computes samples; it only schedules the oscillator's
kernels (through the C ABI, as always) and moves their output.
"""
import torch

from . import _cabi as C
from . import _device as D

__all__ = ["PcmStreamer"]


class PcmStreamer:
    """Iterate over the int16 PCM bytes of `osc`, `chunk_frames` mono frames at a time.

    scale: float->int factor (default 32767; pass 1.0 when the oscillator already carries the
    integer scale, as WaveSynth's *_gen oscillators do).  An out-of-range sample raises
    OverflowError when its chunk is delivered, as upstream's array('h') would.
    max_frames bounds an endless oscillator.  `ahead` chunks are kept in flight."""

    def __init__(self, osc, chunk_frames=1 << 16, scale=32767.0, max_frames=None, ahead=2):
        D.require_cuda()
        assert chunk_frames > 0 and ahead >= 1
        self._osc = osc
        self._chunk = int(chunk_frames)
        self._scale = float(scale)
        self._max = max_frames
        self._ahead = int(ahead)
        self._stream = torch.cuda.Stream()
        n = self._ahead + 1
        self._host = [torch.empty(self._chunk, dtype=torch.int16).pin_memory() for _ in range(n)]
        self._ovf_host = [torch.zeros(1, dtype=torch.int64).pin_memory() for _ in range(n)]
        self._events = [torch.cuda.Event() for _ in range(n)]

    def _launch(self, slot, pos):
        """Render [pos, pos+chunk) on the side stream into ring slot `slot`; returns frames queued."""
        total = self._osc._length()
        end = pos + self._chunk
        if total is not None:
            end = min(end, total)
        if self._max is not None:
            end = min(end, self._max)
        n = end - pos
        if n <= 0:
            return 0
        with torch.cuda.stream(self._stream):
            values = self._osc._fetch(pos, end)
            pcm = D.empty(n, torch.int16)
            ovf = D.zeros(1, torch.int64)
            C.call("sb_f32_to_pcm16", D.ptr(values), n, self._scale, D.ptr(pcm), D.ptr(ovf), D.stream())
            self._host[slot][:n].copy_(pcm, non_blocking=True)
            self._ovf_host[slot].copy_(ovf, non_blocking=True)
            self._events[slot].record(self._stream)
        return n

    def __iter__(self):
        nslots = len(self._host)
        queued = []                    # (slot, frames) in flight, oldest first
        pos = 0
        slot = 0
        done = False
        while True:
            while not done and len(queued) < self._ahead:
                n = self._launch(slot, pos)
                if n == 0:
                    done = True
                    break
                queued.append((slot, n))
                pos += n
                slot = (slot + 1) % nslots
            if not queued:
                return
            s, n = queued.pop(0)
            self._events[s].synchronize()
            bad = int(self._ovf_host[s].item())
            if bad:
                raise OverflowError("%d sample values do not fit 16 bits" % bad)
            yield self._host[s][:n].numpy().tobytes()
