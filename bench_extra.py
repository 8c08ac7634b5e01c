#!/usr/bin/env python
"""
bench_extra.py — the BASELINE.json configurations that are not bench.py's headline, the voice
classes of the bank kernel, the HBM-bound Sample kernels with their roofline fractions, and a
sustained-clock leg.  bench.py calls collect() after its timed region and puts the result into
its JSON line (keys "configs", "classes", "hbm_kernels", "sustained"); run alone it prints one
JSON line per case.

  python bench_extra.py [--reps 5]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import synthesizer_b200 as sb                      # noqa: E402
from synthesizer_b200 import _cabi as C, _device as D   # noqa: E402

SR = 48000


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except (OSError, KeyError):
        return 6650.0, "fallback (B200_PROFILING.md)"


_flush = None


def flush_l2():
    """A 256 MB write (twice the 126 MB L2) so that the next launch finds its inputs in HBM."""
    global _flush
    if _flush is None:
        _flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
    _flush.add_(1)


def timed(fn, reps, flush=True):
    """Median device time (ms, CUDA events on the current stream) of fn over `reps`."""
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if flush:
            flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def wall(fn, reps=3):
    """Median wall-clock ms of fn through the Python API, device synchronised on both sides."""
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


# ------------------------------------------------------------------------------ HBM-bound kernels
def hbm_kernels(reps=5):
    """The integer PCM kernels on a 10-minute stereo 48 kHz buffer (115 MB) and the float filter
    kernels on 10 minutes mono; algorithmic bytes = every operand read once, the result written once."""
    hbm, src = peaks()
    out = {"peak_gbs": hbm, "peak_source": src, "l2": "flushed between repetitions (256 MB write)", "kernels": {}}
    n = 600 * SR * 2
    a = torch.randint(-20000, 20000, (n,), dtype=torch.int16, device="cuda")
    b = torch.randint(-20000, 20000, (n,), dtype=torch.int16, device="cuda")
    o = torch.empty_like(a)
    nin = 44100 * 600
    nout = C.lib.sb_pcm16_ratecv_frames(nin, 44100, 48000)
    dst_t = torch.empty(nout * 2, dtype=torch.int16, device="cuda")
    nf = 600 * SR
    x = torch.rand(nf, device="cuda") * 1.6 - 0.8
    x2 = torch.rand(nf, device="cuda")
    y = torch.empty_like(x)
    p16 = torch.empty(nf, dtype=torch.int16, device="cuda")
    ovf = torch.zeros(1, dtype=torch.int64, device="cuda")
    cases = [
        ("pcm16_add (Sample.mix)", 6 * n, lambda: C.call("sb_pcm16_add", D.ptr(a), D.ptr(b), n, D.ptr(o), D.stream())),
        ("pcm16_mul (Sample.amplify)", 4 * n, lambda: C.call("sb_pcm16_mul", D.ptr(a), n, 0.731, D.ptr(o), D.stream())),
        ("pcm16_fadeout (Sample.fadeout)", 4 * n, lambda: C.call("sb_pcm16_fadeout", D.ptr(o), n, 0.1, D.stream())),
        ("pcm16_tomono (Sample.mono)", 3 * n, lambda: C.call("sb_pcm16_tomono", D.ptr(a), n // 2, 0.5, 0.5, D.ptr(o), D.stream())),
        ("pcm16_ratecv 44100->48000 stereo (Sample.resample)", nin * 4 + nout * 4,
         lambda: C.call("sb_pcm16_ratecv", D.ptr(a), nin, 2, 44100, 48000, D.ptr(dst_t), D.stream())),
        ("f32_to_pcm16", 6 * nf, lambda: C.call("sb_f32_to_pcm16", D.ptr(x), nf, 32767.0, D.ptr(p16), D.ptr(ovf), D.stream())),
        ("mul_f32 (AmpModulationFilter)", 12 * nf, lambda: C.call("sb_mul_f32", D.ptr(x), D.ptr(x2), nf, D.ptr(y), D.stream())),
        ("clip_f32 (ClipFilter)", 8 * nf, lambda: C.call("sb_clip_f32", D.ptr(x), -0.5, 0.5, nf, D.ptr(y), D.stream())),
    ]
    for name, nbytes, fn in cases:
        ms = timed(fn, reps)
        out["kernels"][name] = {"ms": ms, "algorithmic_bytes": nbytes, "gbs": nbytes / ms / 1e6, "frac": nbytes / ms / 1e6 / hbm}
    return out


# ------------------------------------------------------------------------------ configs 1-4
def make_hits(rng, nhits=4096):
    hits = []
    for _ in range(nhits):
        rate = int(rng.choice([22050, 44100, 32000]))
        m = int(rng.integers(2000, 20000))
        t = np.arange(m) / rate
        h = (rng.uniform(8000, 30000) * np.exp(-t * rng.uniform(8, 40)) * np.sin(2 * np.pi * rng.uniform(60, 400) * t)).astype(np.int16)
        hits.append((sb.Sample.from_raw_frames(h.tobytes(), 2, rate, 1), float(rng.uniform(0, 599.0))))
    return hits


def config1(reps=5):
    """BASELINE config 1: a single 440 Hz sine, 1 s mono 44.1 kHz int16 rendered to a Sample."""
    ws = sb.WaveSynth(samplerate=44100)
    fn = lambda: ws.sine(440, 1.0)
    return {"workload": "WaveSynth.sine(440 Hz, 1 s) mono 44.1 kHz int16", "device_ms": timed(fn, reps),
            "wall_ms": wall(fn, 5), "algorithmic_bytes": 88200, "voice_samples": 44100}


def config2(reps=3):
    """BASELINE config 2: Sawtooth FM'd by a Sine, ADSR + echo, 60 s stereo 48 kHz."""
    def fn():
        lfo = sb.Sine(5.0, 0.03, samplerate=SR)
        car = sb.Sawtooth(220.0, 0.45, fm_lfo=lfo, samplerate=SR)
        chain = sb.EchoFilter(sb.EnvelopeFilter(car, 0.5, 1.5, 40.0, 0.6, 10.0), 1.0, 6, 0.35, 0.5)
        return sb.Sample.from_oscillator(chain, 60.0).stereo()
    l0 = D.launch_count()
    s = fn()
    launches = D.launch_count() - l0
    frames = len(s)
    nbytes = frames * 2 * 2
    ms = timed(fn, reps)
    return {"workload": "Sawtooth x Sine FM + ADSR + EchoFilter, 60 s stereo 48 kHz, oscillator API -> Sample",
            "device_ms": ms, "wall_ms": wall(fn), "kernel_launches": launches, "frames": frames,
            "algorithmic_bytes": nbytes, "msamples_per_s": frames / ms / 1e3}


def mixed_bank(nv, rng, amp=None):
    """`nv` voices with carriers and modulators drawn uniformly from the five waveforms (what
    BASELINE config 3 calls 'mixed waveforms'); half of the modulators are sines."""
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    bank = sb.VoiceBank(SR)
    for v in range(nv):
        cw = int(rng.integers(0, 5))
        r = rng.random()
        lfo = None
        if r < 0.5:
            lfo = sb.Sine(float(rng.uniform(0.5, 12)), float(rng.uniform(0.005, 0.06)), samplerate=SR)
        elif r < 0.625:
            lfo = classes[int(rng.integers(1, 5))](float(rng.uniform(0.5, 12)), float(rng.uniform(0.005, 0.06)), samplerate=SR)
        car = classes[cw](float(rng.uniform(55, 1760)), amp or 0.9 / nv, fm_lfo=lfo, samplerate=SR)
        pan = float(rng.uniform(-1, 1))
        bank.add(sb.EnvelopeFilter(car, 0.05, 0.3, 8.0, 0.7, 1.5), (1 - pan) / 2, (1 + pan) / 2)
    return bank


def config3(reps=5):
    """BASELINE config 3: 128-voice chord bank, mixed waveforms, per-voice ADSR, 10 s stereo 48 kHz."""
    bank = mixed_bank(128, np.random.default_rng(3))
    nfr = 10 * SR
    ms = timed(lambda: bank.render(nfr), reps)
    return {"workload": "128-voice chord bank, mixed waveforms + ADSR, 10 s stereo 48 kHz", "device_ms": ms,
            "wall_ms": wall(lambda: bank.to_sample(10.0)), "voice_samples": 128 * nfr, "algorithmic_bytes": nfr * 4,
            "generic_voices": bank.count_generic(), "scan_voices": bank.count_scan(), "msamples_per_s": 128 * nfr / ms / 1e3}


def config4(reps=5, shard=None):
    """BASELINE config 4: mix_at of 4096 resampled drum hits at random offsets into a 10-minute
    48 kHz mono buffer.  Algorithmic bytes: every hit read once and every destination
    span it touches read and written once.  shard=(rank, world): this rank's destination range."""
    hbm, src = peaks()
    rng = np.random.default_rng(0)
    hits = make_hits(rng)
    copies = [h.copy() for h, _ in hits]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sb.Sample.resample_many(copies, SR)
    torch.cuda.synchronize()
    resample_ms = (time.perf_counter() - t0) * 1e3
    placed = [(when, h) for h, (_, when) in zip(copies, hits)]
    base = sb.Sample.from_raw_frames(b"", 2, SR, 1).add_silence(600.0)
    hit_samples = sum(len(h) for _, h in placed)

    def mix_many():
        return base.copy().mix_at_many(placed, shard=shard)

    table = base.placement_table(placed)

    def mix_table():
        return base.copy().mix_at_many(table, shard=shard)

    dst = mix_many()
    torch.cuda.synchronize()
    lo, hi = (dst.shard_range[0], dst.shard_range[1]) if shard else (0, len(dst))
    # the kernel alone on resident data (the same clipping mix_at_many does)
    lens = np.array([len(h) for _, h in placed], dtype=np.int64)
    starts = np.array([int(SR * w) for w, _ in placed], dtype=np.int64)
    pool = torch.cat([h.device_frames() for _, h in placed])
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    o, l, d = sb.Sample.clip_placements(offs, lens, starts, lo, hi)
    idx = torch.tensor(np.stack([o, l, d]), dtype=torch.int64, device="cuda")
    dstbuf = dst.device_frames().clone()
    total = dstbuf.numel()
    nscr = C.lib.sb_pcm16_mix_at_many_scratch_bytes(total, len(l), int(l.sum()))
    scr = torch.empty(nscr, dtype=torch.uint8, device="cuda")
    fn = lambda: C.call("sb_pcm16_mix_at_many", D.ptr(dstbuf), total, D.ptr(pool), D.ptr(idx[0]), D.ptr(idx[1]),
                        D.ptr(idx[2]), len(l), D.ptr(scr), nscr, D.stream())
    ms = timed(fn, reps)
    span = int(l.sum())
    nbytes = 2 * span + 2 * 2 * span
    return {"workload": "Sample.mix_at_many: 4096 resampled drum hits into 10 min @48 kHz mono"
                        + (" (rank %d of %d: destination samples [%d, %d))" % (shard + (lo, hi)) if shard else ""),
            "device_ms": ms, "algorithmic_bytes": nbytes, "gbs": nbytes / ms / 1e6, "hbm_frac": nbytes / ms / 1e6 / hbm,
            "peak_gbs": hbm, "peak_source": src, "hit_samples_in_range": span, "hit_samples": hit_samples,
            "dst_samples": total, "wall_ms": wall(mix_many), "wall_ms_with_table": wall(mix_table),
            "wall_note": "wall_ms: a Python list of 4096 (seconds, Sample) pairs each call; wall_ms_with_table: the same "
                         "hits as a Sample.placement_table built once (both include copying the 57.6 MB destination)",
            "resample_many_wall_ms": resample_ms}


# ------------------------------------------------------------------------------ bank classes
def class_bank(cw, mw, nv=8192):
    """8192 voices of one (carrier, modulator) kind; mw None = no FM, 5 = a Sine of FM gain 3 rad
    (generic kernels); cw 9 = the 25 kinds in equal parts."""
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    b = sb.VoiceBank(SR)
    r = np.random.default_rng(11)
    mixed = cw == 9
    for v in range(nv):
        if mixed:
            cw, mw = v % 5, (v // 5) % 5
        lfo = None
        f0 = float(r.uniform(55, 1760))
        if mw in (5, 6):
            # FM gain 2*pi*f0*depth/sr: depth 16 -> 6.3 - 10.5 rad (generic kernels), depth 3.9 -> 1.5 - 2.6 rad
            # (the fast path's strong-FM variant for Sine / Triangle carriers)
            lfo = sb.Sine(float(r.uniform(0.5, 12)), 16.0 if mw == 5 else 3.9, samplerate=SR)
            f0 = float(r.uniform(3000, 5000))
        elif mw is not None:
            lfo = classes[mw](float(r.uniform(0.5, 12)), float(r.uniform(0.005, 0.06)), samplerate=SR)
        car = classes[cw](f0, 0.9 / nv, fm_lfo=lfo, samplerate=SR)
        pan = float(r.uniform(-1, 1))
        b.add(sb.EnvelopeFilter(car, 0.05, 0.3, 8.0, 0.7, 1.5), (1 - pan) / 2, (1 + pan) / 2)
    return b


CLASSES = [("sine_x_sine", 0, 0, 8192), ("triangle_plain", 1, None, 8192), ("sawtooth_x_sine", 3, 0, 8192),
           ("square_plain", 2, None, 8192), ("sine_x_triangle_lfo", 0, 1, 8192), ("sawtooth_x_square_lfo", 3, 2, 8192),
           ("mixed_5x5", 9, 0, 8192), ("sine_x_sine_strong_fm_2rad", 0, 6, 8192), ("generic_fm_gain_8rad", 0, 5, 1024)]


def classes(reps=3):
    """ms per render of 10 s @ 48 kHz stereo, one voice class at a time (resident voices, exact hints)."""
    nfr = 10 * SR
    out = {}
    for name, cw, mw, nv in CLASSES:
        b = class_bank(cw, mw, nv)
        vd, _ = b._upload(0, None)
        g, sm, e, sc = b.count_classes()
        ms = timed(lambda: b.render(nfr, voices_dev=vd, generic_voices=g, smooth_voices=sm, edged_voices=e, scan_voices=sc), reps, flush=False)
        out[name] = {"voices": nv, "ms": ms, "ms_per_8192_voices": ms * 8192 / nv, "generic_voices": g, "edged_voices": e,
                     "scan_voices": sc, "msamples_per_s": nv * nfr / ms / 1e3}
    return out


# ------------------------------------------------------------------------------ sustained clocks
def sustained(render, seconds=3.0, sampler=None):
    """`render` back to back for at least `seconds` of device time: does the SM clock hold?"""
    render()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    render()
    e1.record()
    torch.cuda.synchronize()
    n = max(10, int(1.1 * seconds * 1e3 / max(e0.elapsed_time(e1), 1e-3)) + 1)      # a tenth over: the first render is the slowest
    if sampler is not None:
        sampler.start()
    e0.record()
    for _ in range(n):
        render()
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop() if sampler is not None else None
    total = e0.elapsed_time(e1)
    return {"renders": n, "seconds": total / 1e3, "ms_per_render": total / n, "clocks": clocks}


def collect(reps=3, only=None):
    out = {}
    parts = {"configs": lambda: {"1": config1(), "2": config2(), "3": config3(), "4": config4()},
             "classes": lambda: classes(reps), "hbm_kernels": lambda: hbm_kernels(max(3, reps))}
    for k, fn in parts.items():
        if only and k not in only:
            continue
        try:
            out[k] = fn()
        except Exception as e:                       # a failing leg must not take the headline line with it
            out[k] = {"error": "%s: %s" % (type(e).__name__, e)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", nargs="*", default=None)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    res = collect(args.reps, args.only)
    for section, cases in res.items():
        if "error" in cases:
            print(json.dumps({"section": section, **cases}), flush=True)
            continue
        body = cases.get("kernels", cases) if section == "hbm_kernels" else cases
        for name, vals in body.items():
            print(json.dumps({"section": section, "case": name, **vals}), flush=True)


if __name__ == "__main__":
    main()
