"""
The cases of the reference-golden comparison, as data.  One table is used three ways:

  tests/golden/make_synthplayer_golden.py   builds each case with the REAL ``synthplayer`` package
                                            (mounted source) and writes its output to an .npz;
  tests/test_reference_golden.py            builds the same cases with the oracle (CPU) and with
                                            synthesizer_b200 (GPU) and compares with that file.

A case is built from a *namespace* ``ns`` that offers the oscillator / filter classes, ``Sample``
and ``WaveSynth`` under upstream's names (``Namespace`` below wraps the three candidates).  Nothing
here depends on which package stands behind it, except white noise: upstream draws from
``random``; the generator records the uniform draws and the other two get them injected
(BASELINE.json north_star: "injecting the same pre-generated noise buffer").

Sizes are small: the whole table renders in seconds in pure Python.  The five BASELINE.json
configurations appear at reduced size under ``config1`` .. ``config5``.
"""
import numpy as np

SR = 48000


class Namespace:
    """Uniform access to one implementation: kind is 'reference', 'oracle' or 'cuda'."""

    def __init__(self, kind, osc_module, sample_class, wavesynth_class, draws=None):
        self.kind = kind
        self.o = osc_module
        self.Sample = sample_class
        self.WaveSynth = wavesynth_class
        self.draws = draws if draws is not None else {}       # case name -> list of uniform draws
        self.recorder = None                                  # reference: set by the generator

    def noise(self, case, frequency, amplitude=1.0, bias=0.0, samplerate=SR):
        if self.kind == "reference":
            self.recorder.begin(case)
            return self.o.WhiteNoise(frequency, amplitude, bias, samplerate=samplerate)
        return self.o.WhiteNoise(frequency, amplitude, bias, samplerate=samplerate,
                                 noise_source=list(self.draws[case]))


# ------------------------------------------------------------------------------ oscillators
# name -> (builder(ns) -> oscillator, frames).  Output: float values of blocks(), flattened.
def _osc_cases():
    c = {}
    waves = ["Sine", "Triangle", "Square", "Sawtooth"]
    for w in waves:
        c["osc_%s_plain" % w] = (lambda ns, w=w: getattr(ns.o, w)(440.0, 0.8, 0.15, 0.05, samplerate=44100), 3000)
        c["osc_%s_divides_rate" % w] = (lambda ns, w=w: getattr(ns.o, w)(1000.0, 0.9, samplerate=SR), 4800)
        c["osc_%s_fm_sine" % w] = (lambda ns, w=w: getattr(ns.o, w)(
            443.7, 0.7, 0.2, 0.0, fm_lfo=ns.o.Sine(5.3, 0.04, 0.1, 0.01, samplerate=SR), samplerate=SR), 6000)
        c["osc_%s_fm_triangle" % w] = (lambda ns, w=w: getattr(ns.o, w)(
            330.0, 0.6, fm_lfo=ns.o.Triangle(3.0, 0.05, samplerate=SR), samplerate=SR), 6000)
    c["osc_Pulse_plain"] = (lambda ns: ns.o.Pulse(440.0, 0.8, 0.15, 0.05, 0.3, samplerate=44100), 3000)
    c["osc_Pulse_fm_sine"] = (lambda ns: ns.o.Pulse(
        443.7, 0.7, 0.2, 0.0, 0.3, fm_lfo=ns.o.Sine(5.3, 0.04, 0.1, 0.01, samplerate=SR), samplerate=SR), 6000)
    c["osc_Pulse_pwm"] = (lambda ns: ns.o.Pulse(
        220.0, 0.7, pulsewidth=0.4, pwm_lfo=ns.o.Sine(2.0, 0.3, bias=0.5, samplerate=SR), samplerate=SR), 6000)
    c["osc_Sine_fm_square"] = (lambda ns: ns.o.Sine(
        660.0, 0.9, fm_lfo=ns.o.Square(7.0, 0.02, samplerate=SR), samplerate=SR), 6000)
    c["osc_Sawtooth_fm_pulse"] = (lambda ns: ns.o.Sawtooth(
        220.0, 0.9, fm_lfo=ns.o.Pulse(4.0, 0.03, pulsewidth=0.25, samplerate=SR), samplerate=SR), 6000)
    c["osc_Sine_fm_nested"] = (lambda ns: ns.o.Sine(330.0, 0.8, 0.1, fm_lfo=ns.o.Sine(
        6.0, 0.04, 0.2, fm_lfo=ns.o.Sine(0.7, 0.3, samplerate=SR), samplerate=SR), samplerate=SR), 6000)
    c["osc_Harmonics"] = (lambda ns: ns.o.Harmonics(220.0, [(1, 1.0), (2, 0.5), (5, 0.2)], 0.8, samplerate=SR), 3000)
    c["osc_SquareH"] = (lambda ns: ns.o.SquareH(220.0, 7, 0.8, samplerate=SR), 3000)
    c["osc_SawtoothH"] = (lambda ns: ns.o.SawtoothH(220.0, 7, 0.8, samplerate=SR), 3000)
    c["osc_Semicircle"] = (lambda ns: ns.o.Semicircle(330.0, 0.8, 0.1, 0.05, samplerate=SR), 3000)
    c["osc_Pointy"] = (lambda ns: ns.o.Pointy(330.0, 0.8, 0.1, 0.05, samplerate=SR), 3000)
    c["osc_Linear"] = (lambda ns: ns.o.Linear(0.2, 0.001, -0.5, 0.9, samplerate=SR), 3000)
    c["osc_Linear_start_outside"] = (lambda ns: ns.o.Linear(1.5, -0.002, -1.0, 1.0, samplerate=SR), 1500)
    for w in ["FastSine", "FastTriangle", "FastSquare", "FastSawtooth"]:
        c["osc_%s" % w] = (lambda ns, w=w: getattr(ns.o, w)(440.0, 0.8, 0.15, 0.05, samplerate=44100), 3000)
    c["osc_FastPulse"] = (lambda ns: ns.o.FastPulse(440.0, 0.8, 0.15, 0.05, 0.3, samplerate=44100), 3000)
    c["osc_WhiteNoise"] = (lambda ns: ns.noise("osc_WhiteNoise", 4000.0, 0.7, 0.05), 3000)
    # ---- filters
    fm = lambda ns: ns.o.Sawtooth(220.0, 0.45, fm_lfo=ns.o.Sine(5.0, 0.03, samplerate=SR), samplerate=SR)
    c["flt_Envelope"] = (lambda ns: ns.o.EnvelopeFilter(ns.o.Sine(440.0, 0.9, samplerate=SR), 0.01, 0.02, 0.03, 0.6, 0.04), 6000)
    c["flt_Envelope_stop_at_end"] = (lambda ns: ns.o.EnvelopeFilter(
        ns.o.Sine(440.0, 0.9, samplerate=SR), 0.01, 0.02, 0.03, 0.6, 0.04, stop_at_end=True), 6000)
    c["flt_Envelope_exact_boundaries"] = (lambda ns: ns.o.EnvelopeFilter(
        ns.o.Triangle(100.0, 0.9, samplerate=SR), 0.0625, 0.0625, 0.0, 0.5, 0.125), 14000)
    c["flt_AmpModulation"] = (lambda ns: ns.o.AmpModulationFilter(
        ns.o.Sine(440.0, 0.9, samplerate=SR), ns.o.Triangle(6.0, 0.5, bias=0.5, samplerate=SR)), 6000)
    c["flt_Mixing"] = (lambda ns: ns.o.MixingFilter(
        ns.o.Sine(440.0, 0.3, samplerate=SR), ns.o.Square(110.0, 0.2, samplerate=SR), fm(ns)), 6000)
    c["flt_Delay"] = (lambda ns: ns.o.DelayFilter(ns.o.Sine(440.0, 0.9, samplerate=SR), 0.0123), 3000)
    c["flt_Delay_negative"] = (lambda ns: ns.o.DelayFilter(ns.o.Sine(440.0, 0.9, samplerate=SR), -0.0123), 3000)
    c["flt_Echo"] = (lambda ns: ns.o.EchoFilter(
        ns.o.EnvelopeFilter(ns.o.Sine(440.0, 0.5, samplerate=SR), 0.005, 0.005, 0.01, 0.5, 0.01, stop_at_end=True),
        0.02, 3, 0.03, 0.5), 12000)
    c["flt_Clip"] = (lambda ns: ns.o.ClipFilter(ns.o.Sine(440.0, 0.9, samplerate=SR), -0.3, 0.5), 3000)
    c["flt_Abs"] = (lambda ns: ns.o.AbsFilter(ns.o.Sawtooth(440.0, 0.9, samplerate=SR)), 3000)
    c["flt_Null"] = (lambda ns: ns.o.NullFilter(ns.o.Sine(440.0, 0.9, samplerate=SR)), 1000)
    # ---- BASELINE.json configurations 2, 3, 5 at reduced size (float stream; PCM under SAMPLE_CASES)
    c["config2_chain"] = (lambda ns: ns.o.EchoFilter(ns.o.EnvelopeFilter(fm(ns), 0.02, 0.05, 0.1, 0.6, 0.08,
                                                                         stop_at_end=True), 0.05, 4, 0.06, 0.5), 24000)
    c["config3_bank"] = (lambda ns: ns.o.MixingFilter(*bank_voices(ns, 16, seed=3, mixed=True)), 12000)
    c["config5_bank"] = (lambda ns: ns.o.MixingFilter(*bank_voices(ns, 24, seed=5, mixed=False)), 12000)
    return c


def bank_voices(ns, nvoices, seed, mixed):
    """The voices of a reduced config 3 (mixed waveforms) or config 5 (Sine x Sine) bank: FM
    carrier inside an ADSR envelope.  All parameters are float32-representable so that a
    C-ABI sb_voice holds them exactly."""
    rng = np.random.default_rng(seed)
    f32 = lambda x: float(np.float32(x))
    waves = ["Sine", "Triangle", "Square", "Sawtooth", "Pulse"]
    voices = []
    for v in range(nvoices):
        cw = waves[int(rng.integers(0, 5))] if mixed else "Sine"
        mw = waves[int(rng.integers(0, 5))] if (mixed and rng.random() < 0.3) else "Sine"
        lfo = getattr(ns.o, mw)(f32(rng.uniform(0.5, 12)), f32(rng.uniform(0.005, 0.05)), samplerate=SR)
        car = getattr(ns.o, cw)(f32(rng.uniform(55, 1760)), f32(0.9 / nvoices), f32(rng.random()), fm_lfo=lfo,
                                samplerate=SR)
        voices.append(ns.o.EnvelopeFilter(car, f32(rng.uniform(0.005, 0.03)), f32(rng.uniform(0.01, 0.05)),
                                          f32(rng.uniform(0.05, 0.1)), f32(rng.uniform(0.4, 0.9)),
                                          f32(rng.uniform(0.02, 0.08))))
    return voices


OSC_CASES = _osc_cases()


# ------------------------------------------------------------------------------ Sample
def _pcm(seed, n, lo=-20000, hi=20000):
    return np.random.default_rng(seed).integers(lo, hi, n).astype("<i2")


def _s(ns, seed, frames, channels=1, rate=SR, lo=-20000, hi=20000):
    return ns.Sample.from_raw_frames(_pcm(seed, frames * channels, lo, hi).tobytes(), 2, rate, channels)


def _s24(ns, seed, frames, channels=1, rate=SR):
    v = np.random.default_rng(seed).integers(-(1 << 23), 1 << 23, frames * channels).astype(np.int64) & 0xFFFFFF
    raw = np.stack([v & 0xFF, (v >> 8) & 0xFF, (v >> 16) & 0xFF], axis=1).astype(np.uint8).tobytes()
    return ns.Sample.from_raw_frames(raw, 3, rate, channels)


# name -> builder(ns) -> Sample.  Output: the sample's frames (bytes) + rate, width, channels.
def _sample_cases():
    c = {}
    c["smp_mix"] = lambda ns: _s(ns, 1, 4000).mix(_s(ns, 2, 3000, lo=-32768, hi=32767))
    c["smp_mix_at"] = lambda ns: _s(ns, 1, 4000).mix_at(0.03, _s(ns, 2, 3000, lo=-32768, hi=32767))
    c["smp_mix_at_extends"] = lambda ns: _s(ns, 1, 4000).mix_at(0.07, _s(ns, 2, 3000))
    c["smp_mix_stereo"] = lambda ns: _s(ns, 1, 4000, 2).mix(_s(ns, 2, 4000, 2, lo=-32768, hi=32767))
    c["smp_amplify"] = lambda ns: _s(ns, 3, 4000, 2).amplify(1.731)
    c["smp_amplify_max"] = lambda ns: _s(ns, 3, 4000, 2, lo=-9000, hi=7000).amplify_max()
    c["smp_at_volume"] = lambda ns: _s(ns, 3, 4000).at_volume(0.37)
    c["smp_fadein"] = lambda ns: _s(ns, 4, 4800).fadein(0.05, 0.1)
    c["smp_fadeout"] = lambda ns: _s(ns, 4, 4800).fadeout(0.05, 0.2)
    c["smp_fadeout_stereo"] = lambda ns: _s(ns, 4, 4800, 2).fadeout(0.03)
    c["smp_envelope"] = lambda ns: _s(ns, 5, 9600).envelope(0.02, 0.03, 0.6, 0.04)
    c["smp_echo"] = lambda ns: _s(ns, 6, 2400, lo=-8000, hi=8000).echo(0.05, 3, 0.02, 0.6)
    c["smp_resample_up"] = lambda ns: _s(ns, 7, 4410, 1, 44100).resample(48000)
    c["smp_resample_down_stereo"] = lambda ns: _s(ns, 7, 4800, 2, 48000).resample(22050)
    c["smp_speed"] = lambda ns: _s(ns, 7, 4800).speed(1.5)
    c["smp_mono"] = lambda ns: _s(ns, 8, 3000, 2).mono(0.8, 0.6)
    c["smp_stereo"] = lambda ns: _s(ns, 8, 3000, 1).stereo(0.8, 0.6)
    c["smp_left"] = lambda ns: _s(ns, 8, 3000, 2).left()
    c["smp_right"] = lambda ns: _s(ns, 8, 3000, 2).right()
    c["smp_pan"] = lambda ns: _s(ns, 8, 3000, 1).pan(0.3)
    c["smp_invert"] = lambda ns: _s(ns, 9, 3000, lo=-32768, hi=32767).invert()
    c["smp_bias"] = lambda ns: _s(ns, 9, 3000).bias(5000)
    c["smp_reverse"] = lambda ns: _s(ns, 9, 3000, 2).reverse()
    c["smp_normalize_from_8k_mono"] = lambda ns: _s(ns, 9, 800, 1, 8000).normalize()
    c["smp_make_32bit"] = lambda ns: _s(ns, 9, 3000).make_32bit()
    c["smp_make_16bit_from_32"] = lambda ns: _s(ns, 9, 3000, lo=-9000, hi=7000).make_32bit().make_16bit()
    c["smp_clip"] = lambda ns: _s(ns, 10, 4800).clip(0.01, 0.05)
    c["smp_split"] = lambda ns: _s(ns, 10, 4800).split(0.04)
    c["smp_add_silence"] = lambda ns: _s(ns, 10, 1000).add_silence(0.01).add_silence(0.005, at_start=True)
    c["smp_join"] = lambda ns: _s(ns, 10, 1000).join(_s(ns, 11, 700))
    c["smp_delay"] = lambda ns: _s(ns, 10, 2000).delay(0.01)
    c["smp_delay_keep_length"] = lambda ns: _s(ns, 10, 2000).delay(0.01, keep_length=True)
    c["smp_delay_negative"] = lambda ns: _s(ns, 10, 2000).delay(-0.01)
    c["smp_modulate_amp_osc"] = lambda ns: _s(ns, 12, 4800).modulate_amp(ns.o.Sine(20.0, 0.5, bias=0.5, samplerate=SR))
    c["smp_modulate_amp_sample"] = lambda ns: _s(ns, 12, 4800).modulate_amp(
        ns.Sample.from_raw_frames(np.array([0, 30, 70, 100, -70], dtype="<i2").tobytes(), 2, SR, 1))
    c["smp_stereo_mix"] = lambda ns: _s(ns, 13, 3000, 1).stereo_mix(_s(ns, 14, 2000, 1), "R", 0.7, 0.01)
    # 24-bit frames (audioop size 3)
    c["smp24_amplify"] = lambda ns: _s24(ns, 15, 3000).amplify(1.731)
    c["smp24_mix_at"] = lambda ns: _s24(ns, 15, 4000).mix_at(0.03, _s24(ns, 16, 3000))
    c["smp24_resample_stereo"] = lambda ns: _s24(ns, 15, 2400, 2).resample(44100)
    c["smp24_make_16bit"] = lambda ns: _s24(ns, 15, 3000).make_16bit()
    # float -> PCM (K1), through both entry points
    c["smp_from_oscillator"] = lambda ns: ns.Sample.from_oscillator(ns.o.Sine(440.0, 0.9, samplerate=SR), 0.05)
    c["smp_from_osc_block"] = lambda ns: ns.Sample.from_osc_block(
        next(iter(ns.o.Sawtooth(440.0, 0.9, samplerate=SR).blocks())), SR)
    # BASELINE.json config 1: a single 440 Hz sine, 1 s mono 44.1 kHz int16, through WaveSynth
    c["config1_wavesynth_sine"] = lambda ns: ns.WaveSynth(samplerate=44100).sine(440, 1.0)
    c["wavesynth_square_fm"] = lambda ns: ns.WaveSynth(samplerate=SR).square(
        220, 0.1, fm_lfo=ns.o.Sine(5.0, 0.03, samplerate=SR))
    c["wavesynth_pulse"] = lambda ns: ns.WaveSynth(samplerate=SR).pulse(220, 0.1, pulsewidth=0.3)
    c["wavesynth_triangle_32bit"] = lambda ns: ns.WaveSynth(samplerate=SR, samplewidth=4).triangle(220, 0.05)
    c["wavesynth_harmonics"] = lambda ns: ns.WaveSynth(samplerate=SR).harmonics(220, 0.05, [(1, 1.0), (3, 0.4), (4, 0.2)], amplitude=0.5)
    c["wavesynth_square_h"] = lambda ns: ns.WaveSynth(samplerate=SR).square_h(220, 0.05, 5, amplitude=0.6)
    c["wavesynth_sawtooth_h"] = lambda ns: ns.WaveSynth(samplerate=SR).sawtooth_h(220, 0.05, 5, amplitude=0.4)
    c["wavesynth_semicircle"] = lambda ns: ns.WaveSynth(samplerate=SR).semicircle(330, 0.05)
    c["wavesynth_pointy"] = lambda ns: ns.WaveSynth(samplerate=SR).pointy(330, 0.05, amplitude=0.8)
    # BASELINE.json config 2 at reduced size: the chain rendered to a Sample and made stereo
    c["config2_sample"] = lambda ns: ns.Sample.from_oscillator(OSC_CASES["config2_chain"][0](ns), 0.5).stereo()
    # BASELINE.json config 4 at reduced size: resampled hits mixed at offsets, in list order
    c["config4_mix_at"] = _config4
    return c


def _config4(ns):
    rng = np.random.default_rng(4)
    dst = ns.Sample.from_raw_frames(b"", 2, SR, 1).add_silence(1.0)
    for k in range(24):
        rate = int(rng.choice([22050, 44100, 32000]))
        m = int(rng.integers(200, 2000))
        t = np.arange(m) / rate
        h = (rng.uniform(8000, 30000) * np.exp(-t * rng.uniform(8, 40)) *
             np.sin(2 * np.pi * rng.uniform(60, 400) * t)).astype("<i2")
        hit = ns.Sample.from_raw_frames(h.tobytes(), 2, rate, 1).resample(SR)
        dst.mix_at(float(np.float32(rng.uniform(0, 0.95))), hit)
    return dst


SAMPLE_CASES = _sample_cases()


# ------------------------------------------------------------------------------ running a case
def render_osc(ns, name):
    """float64 array of the first `frames` values of the case's blocks() (shorter if it ends)."""
    build, frames = OSC_CASES[name]
    osc = build(ns)
    if ns.kind == "cuda":
        # the GPU backend renders a range at once; blocks() is the same values in 512-float lists
        n = frames
        length = osc._length() if hasattr(osc, "_length") else None
        if length is not None:
            n = min(n, length)
        return osc.render(n).double().cpu().numpy()
    out = []
    for block in osc.blocks():
        out.extend(block)
        if len(out) >= frames:
            break
    return np.asarray(out[:frames], dtype=np.float64)


def render_sample(ns, name):
    """(int array of sample values, samplerate, samplewidth, nchannels) of the case's Sample."""
    s = SAMPLE_CASES[name](ns)
    if hasattr(s, "get_frames"):
        raw = s.get_frames()
    else:
        raw = bytes(s.view_frame_data())
    width = s.samplewidth
    if width == 3:
        b = np.frombuffer(raw, dtype=np.uint8).reshape(-1, 3).astype(np.int32)
        vals = (b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16))
        vals = np.where(vals & 0x800000, vals - 0x1000000, vals)
    else:
        vals = np.frombuffer(raw, dtype={2: "<i2", 4: "<i4"}[width]).astype(np.int64)
    return vals.astype(np.int64), int(s.samplerate), int(width), int(s.nchannels)
