#!/usr/bin/env python
"""
bench_extra.py — the BASELINE.json configurations that are not bench.py's headline line, and
the HBM-bound integer kernels with their roofline fractions.  One JSON line per case.
Not part of the driver's contract; results are kept under profiles/.

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


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except (OSError, KeyError):
        return 6650.0, "fallback"


def flush_l2(buf):
    buf.add_(1)


def timed(fn, reps, flush=None):
    """Median device time (ms) of fn over `reps`, L2 flushed (a 512 MB write) between runs."""
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush_l2(flush)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def line(name, **kw):
    print(json.dumps({"case": name, **kw}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    hbm, src = peaks()
    flush = torch.zeros(512 << 20, dtype=torch.uint8, device="cuda")
    rng = np.random.default_rng(0)

    # ---- HBM-bound PCM kernels on a 10-minute stereo 48 kHz buffer (115 MB) -------------------
    n = 600 * 48000 * 2
    a = torch.randint(-20000, 20000, (n,), dtype=torch.int16, device="cuda")
    b = torch.randint(-20000, 20000, (n,), dtype=torch.int16, device="cuda")
    out = torch.empty_like(a)
    cases = [
        ("pcm16_add (Sample.mix)", 6 * n, lambda: C.call("sb_pcm16_add", D.ptr(a), D.ptr(b), n, D.ptr(out), D.stream())),
        ("pcm16_mul (Sample.amplify)", 4 * n, lambda: C.call("sb_pcm16_mul", D.ptr(a), n, 0.731, D.ptr(out), D.stream())),
        ("pcm16_fadeout (Sample.fadeout)", 4 * n, lambda: C.call("sb_pcm16_fadeout", D.ptr(out), n, 0.1, D.stream())),
        ("pcm16_tomono (Sample.mono)", 3 * n, lambda: C.call("sb_pcm16_tomono", D.ptr(a), n // 2, 0.5, 0.5, D.ptr(out), D.stream())),
        ("pcm16_ratecv 44100->48000 stereo (Sample.resample)", None, None),
    ]
    for name, nbytes, fn in cases:
        if fn is None:
            nin = 44100 * 600
            src_t = a[: nin * 2]
            nout = C.lib.sb_pcm16_ratecv_frames(nin, 44100, 48000)
            dst_t = torch.empty(nout * 2, dtype=torch.int16, device="cuda")
            fn = lambda: C.call("sb_pcm16_ratecv", D.ptr(src_t), nin, 2, 44100, 48000, D.ptr(dst_t), D.stream())
            nbytes = nin * 4 + nout * 4
        ms = timed(fn, args.reps, flush)
        gbs = nbytes / ms / 1e6
        line(name, ms=ms, algorithmic_bytes=nbytes, achieved_gbs=gbs, peak_gbs=hbm, peak_source=src, frac=gbs / hbm)

    # float -> PCM and the float filter kernels (4 B in, 2 or 4 B out)
    nf = 600 * 48000
    x = torch.rand(nf, device="cuda") * 1.6 - 0.8
    x2 = torch.rand(nf, device="cuda")
    y = torch.empty_like(x)
    p16 = torch.empty(nf, dtype=torch.int16, device="cuda")
    ovf = torch.zeros(1, dtype=torch.int64, device="cuda")
    for name, nbytes, fn in [
        ("f32_to_pcm16", 6 * nf, lambda: C.call("sb_f32_to_pcm16", D.ptr(x), nf, 32767.0, D.ptr(p16), D.ptr(ovf), D.stream())),
        ("mul_f32 (AmpModulationFilter)", 12 * nf, lambda: C.call("sb_mul_f32", D.ptr(x), D.ptr(x2), nf, D.ptr(y), D.stream())),
        ("clip_f32 (ClipFilter)", 8 * nf, lambda: C.call("sb_clip_f32", D.ptr(x), -0.5, 0.5, nf, D.ptr(y), D.stream())),
    ]:
        ms = timed(fn, args.reps, flush)
        gbs = nbytes / ms / 1e6
        line(name, ms=ms, algorithmic_bytes=nbytes, achieved_gbs=gbs, peak_gbs=hbm, peak_source=src, frac=gbs / hbm)

    # ---- config 4: mix_at of 4096 resampled drum hits into a 10-minute 48 kHz mono buffer -------
    sr = 48000
    hits = []
    for k in range(4096):
        src_rate = int(rng.choice([22050, 44100, 32000]))
        m = int(rng.integers(2000, 20000))
        t = np.arange(m) / src_rate
        h = (rng.uniform(8000, 30000) * np.exp(-t * rng.uniform(8, 40)) * np.sin(2 * np.pi * rng.uniform(60, 400) * t)).astype(np.int16)
        hits.append((src_rate, sb.Sample.from_raw_frames(h.tobytes(), 2, src_rate, 1), float(rng.uniform(0, 599.0))))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    placed = [(when, h.copy().resample(sr)) for _, h, when in hits]
    torch.cuda.synchronize()
    resample_ms = (time.perf_counter() - t0) * 1e3
    copies = [h.copy() for _, h, _ in hits]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sb.Sample.resample_many(copies, sr)
    torch.cuda.synchronize()
    resample_many_ms = (time.perf_counter() - t0) * 1e3
    base = sb.Sample.from_raw_frames(b"", 2, sr, 1).add_silence(600.0)
    span_samples = sum(len(h) for _, h in placed)

    def mix_many():
        dst = base.copy()
        dst.mix_at_many(placed)
        return dst

    # kernel-only timing of sb_pcm16_mix_at_many (device-resident pool and indices)
    lens = [len(h) for _, h in placed]
    starts = [int(sr * w) for w, _ in placed]
    pool = torch.cat([h.device_frames() for _, h in placed])
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]])
    idx = torch.tensor(np.stack([offs, lens, starts]), dtype=torch.int64, device="cuda")
    dstbuf = base.copy().device_frames()
    total = dstbuf.numel()
    nscr = C.lib.sb_pcm16_mix_at_many_scratch_bytes(total, len(lens), int(sum(lens)))
    scr = torch.empty(nscr, dtype=torch.uint8, device="cuda")
    fn = lambda: C.call("sb_pcm16_mix_at_many", D.ptr(dstbuf), total, D.ptr(pool), D.ptr(idx[0]), D.ptr(idx[1]),
                        D.ptr(idx[2]), len(lens), D.ptr(scr), nscr, D.stream())
    ms = timed(fn, args.reps, flush)
    fn0 = lambda: C.call("sb_pcm16_mix_at_many", D.ptr(dstbuf), total, D.ptr(pool), D.ptr(idx[0]), D.ptr(idx[1]),
                         D.ptr(idx[2]), len(lens), None, 0, D.stream())
    ms_noindex = timed(fn0, args.reps, flush)
    # algorithmic bytes: every source read once, every touched destination span read and written once
    nbytes = 2 * span_samples + 2 * 2 * span_samples
    line("config4 mix_at_many kernel: 4096 hits into 10 min @48 kHz mono", ms=ms, algorithmic_bytes=nbytes,
         achieved_gbs=nbytes / ms / 1e6, peak_gbs=hbm, peak_source=src, frac=nbytes / ms / 1e6 / hbm,
         hit_samples=span_samples, dst_samples=total, ms_without_index=ms_noindex)
    t0 = time.perf_counter()
    mix_many()
    torch.cuda.synchronize()
    api_ms = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    dst = base.copy()
    for when, h in placed[:512]:
        dst.mix_at(when, h)
    torch.cuda.synchronize()
    loop_ms = (time.perf_counter() - t0) * 1e3 * 8
    line("config4 through the Sample API", resample_4096_hits_ms=resample_ms, resample_many_ms=resample_many_ms,
         mix_at_many_ms=api_ms,
         mix_at_loop_ms_extrapolated_from_512=loop_ms, msamples_per_s=total / api_ms / 1e3)

    # ---- config 2: FM sawtooth + ADSR + echo, 60 s stereo 48 kHz through the oscillator API -------
    def config2():
        lfo = sb.Sine(5.0, 0.03, samplerate=sr)
        car = sb.Sawtooth(220.0, 0.45, fm_lfo=lfo, samplerate=sr)
        chain = sb.EchoFilter(sb.EnvelopeFilter(car, 0.5, 1.5, 40.0, 0.6, 10.0), 1.0, 6, 0.35, 0.5)
        return sb.Sample.from_oscillator(chain, 60.0).stereo()
    config2()
    torch.cuda.synchronize()
    l0 = D.launch_count()
    t0 = time.perf_counter()
    s = config2()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    line("config2 FM saw + ADSR + echo, 60 s stereo 48 kHz (oscillator API -> Sample)", wall_ms=wall,
         kernel_launches=D.launch_count() - l0, frames=len(s), msamples_per_s=2 * len(s) / wall / 1e3)

    # ---- config 1: 440 Hz sine, 1 s mono 44.1 kHz -> Sample (WaveSynth) -----------------------------
    ws = sb.WaveSynth(samplerate=44100)
    ws.sine(440, 1.0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        ws.sine(440, 1.0)
    torch.cuda.synchronize()
    line("config1 WaveSynth.sine(440, 1 s) mono 44.1 kHz", wall_ms=(time.perf_counter() - t0) * 1e3 / 20)

    # ---- config 3: 128-voice chord bank (mixed waveforms), 10 s stereo 48 kHz ----------------------
    classes = [sb.Sine, sb.Triangle, sb.Square, sb.Sawtooth, sb.Pulse]
    bank = sb.VoiceBank(sr)
    for v in range(128):
        # mixed carrier waveforms; half of the voices have a sine vibrato, an eighth an LFO of
        # another waveform (those take the tile-scan variant of the fast kernel)
        cw = int(rng.integers(0, 5))
        r = rng.random()
        lfo = None
        if r < 0.5:
            lfo = sb.Sine(float(rng.uniform(0.5, 12)), float(rng.uniform(0.005, 0.06)), samplerate=sr)
        elif r < 0.625:
            lfo = classes[int(rng.integers(1, 5))](float(rng.uniform(0.5, 12)), float(rng.uniform(0.005, 0.06)), samplerate=sr)
        car = classes[cw](float(rng.uniform(55, 1760)), 0.9 / 128, fm_lfo=lfo, samplerate=sr)
        pan = float(rng.uniform(-1, 1))
        bank.add(sb.EnvelopeFilter(car, 0.05, 0.3, 8.0, 0.7, 1.5), (1 - pan) / 2, (1 + pan) / 2)
    nfr = 10 * sr
    ms = timed(lambda: bank.render(nfr), args.reps)
    line("config3 128-voice chord bank, mixed waveforms + ADSR, 10 s stereo 48 kHz", ms=ms,
         generic_voices=bank.count_generic(), msamples_per_s=128 * nfr / ms / 1e3)

    # ---- one kernel class at a time: 8192 voices x 10 s of a single (carrier, modulator) kind -------
    def class_bank(cw, mw, nv=8192):
        b = sb.VoiceBank(sr)
        r = np.random.default_rng(11)
        for v in range(nv):
            lfo = None
            f0 = float(r.uniform(55, 1760))
            if mw == 5:          # a Sine modulator too strong for the fast path: 3.9 x 2 pi f0 / sr > 1.5 rad
                lfo = sb.Sine(float(r.uniform(0.5, 12)), 3.9, samplerate=sr)
                f0 = float(r.uniform(3000, 5000))
            elif mw is not None:
                lfo = classes[mw](float(r.uniform(0.5, 12)), float(r.uniform(0.005, 0.06)), samplerate=sr)
            car = classes[cw](f0, 0.9 / nv, fm_lfo=lfo, samplerate=sr)
            pan = float(r.uniform(-1, 1))
            b.add(sb.EnvelopeFilter(car, 0.05, 0.3, 8.0, 0.7, 1.5), (1 - pan) / 2, (1 + pan) / 2)
        return b
    for name, cw, mw, nv in [("Sine x Sine (fast, smooth)", 0, 0, 8192), ("Triangle, no FM (fast, smooth)", 1, None, 8192),
                             ("Sawtooth x Sine (fast, edged)", 3, 0, 8192), ("Square, no FM (fast, edged)", 2, None, 8192),
                             ("Sine x Triangle LFO (fast, tile scan)", 0, 1, 8192), ("Sawtooth x Square LFO (fast, tile scan)", 3, 2, 8192),
                             ("Sine x Sine, FM gain 3 rad (generic)", 0, 5, 1024)]:
        b = class_bank(cw, mw, nv)
        vd, _ = b._upload(0, None)                       # voice parameters resident, as in bench.py
        g, e, sc = b.count_generic(), b.count_edged(), b.count_scan()
        ms = timed(lambda: b.render(nfr, voices_dev=vd, generic_voices=g, edged_voices=e, scan_voices=sc),
                   max(2, args.reps // 4))
        line("bank class: %s, %d voices x 10 s" % (name, nv), ms=ms, generic_voices=g, edged_voices=e, scan_voices=sc,
             msamples_per_s=nv * nfr / ms / 1e3)


if __name__ == "__main__":
    main()
