"""
GPU parity of Sample against the oracle Sample (oracle/sample_oracle.py), whose integer
arithmetic is CPython's stdlib audioop executed directly.  Integer PCM: bit exact.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pair(frames, sr=44100, ch=1):
    import synthesizer_b200 as sb
    from oracle.sample_oracle import Sample as OS
    raw = np.asarray(frames, dtype="<i2").tobytes()
    return sb.Sample.from_raw_frames(raw, 2, sr, ch), OS.from_raw_frames(raw, 2, sr, ch)


def _same(g, o, what=""):
    assert (g.samplerate, g.nchannels, g.samplewidth) == (o.samplerate, o.nchannels, o.samplewidth), what
    a = np.frombuffer(g.get_frames(), dtype="<i2")
    b = np.frombuffer(o._frames, dtype="<i2")
    assert a.shape == b.shape, "%s: %s vs %s" % (what, a.shape, b.shape)
    if not np.array_equal(a, b):
        i = int(np.nonzero(a != b)[0][0])
        raise AssertionError("%s: first difference at %d: got %d want %d (%d differ)"
                             % (what, i, a[i], b[i], int((a != b).sum())))


def _rand(n, seed, full=True):
    rng = np.random.default_rng(seed)
    x = rng.integers(-32768, 32768, size=n) if full else rng.integers(-12000, 12000, size=n)
    if n > 8:
        x[:4] = [32767, -32768, -32767, 0]
    return x.astype(np.int16)


def test_construct_and_roundtrip(cuda):
    g, o = _pair(_rand(1001, 1))
    _same(g, o)
    assert len(g) == len(o) and g.duration == o.duration and g.maximum == o.maximum
    e, eo = _pair([])
    assert len(e) == 0 and e.get_frames() == b"" and e.duration == 0


@pytest.mark.parametrize("factor", [0.5, 1.5, -1.0, 0.333, 2.75, 0.0, -0.9999, 1e-3])
def test_amplify(cuda, factor):
    g, o = _pair(_rand(50001, 2))
    _same(g.amplify(factor), o.amplify(factor), "amplify %r" % factor)


def test_amplify_max_invert_bias_reverse(cuda):
    g, o = _pair(_rand(4097, 3, full=False))
    _same(g.amplify_max(), o.amplify_max(), "amplify_max")
    _same(g.invert(), o.invert(), "invert")
    _same(g.bias(1234), o.bias(1234), "bias")
    _same(g.bias(-30000), o.bias(-30000), "bias wrap")
    _same(g.reverse(), o.reverse(), "reverse")


@pytest.mark.parametrize("n1,n2", [(1000, 1000), (1000, 3001), (3001, 1000), (7, 0)])
def test_mix(cuda, n1, n2):
    g1, o1 = _pair(_rand(n1, 4))
    g2, o2 = _pair(_rand(n2, 5))
    _same(g1.mix(g2), o1.mix(o2), "mix")
    g1, o1 = _pair(_rand(5000, 6))
    g2, o2 = _pair(_rand(44100, 7))
    _same(g1.mix(g2, other_seconds=0.05), o1.mix(o2, other_seconds=0.05), "mix other_seconds")


@pytest.mark.parametrize("ch", [1, 2])
def test_mix_at(cuda, ch):
    g1, o1 = _pair(_rand(44100 * ch, 8), ch=ch)
    g2, o2 = _pair(_rand(10001 * ch, 9), ch=ch)
    for t in (0.0, 0.1234, 0.5, 0.95, 1.7):
        g1.mix_at(t, g2)
        o1.mix_at(t, o2)
        _same(g1, o1, "mix_at %g" % t)
    g1.mix_at(0.3, g2, other_seconds=0.1)
    o1.mix_at(0.3, o2, other_seconds=0.1)
    _same(g1, o1, "mix_at other_seconds")


def test_mix_at_many_equals_sequential_with_saturation(cuda):
    import synthesizer_b200 as sb
    rng = np.random.default_rng(10)
    sr = 8000
    g, o = _pair(_rand(sr * 3, 11), sr=sr)
    placements_g, placements_o = [], []
    for k in range(300):
        n = int(rng.integers(1, 3000))
        t = float(rng.uniform(0, 3.2))
        hg, ho = _pair(_rand(n, 100 + k), sr=sr)       # full scale: saturates often
        placements_g.append((t, hg))
        placements_o.append((t, ho))
    for t, h in placements_o:
        o.mix_at(t, h)
    g.mix_at_many(placements_g)
    _same(g, o, "mix_at_many")


@pytest.mark.parametrize("rates", [(44100, 48000), (48000, 44100), (22050, 48000), (48000, 8000),
                                   (44100, 44101), (8000, 96000), (11025, 44100), (96000, 44100),
                                   (1999999, 2000000), (2000000, 1999999), (48000, 16), (16, 1600),
                                   (65537, 65536), (1048573, 1048576)])
@pytest.mark.parametrize("ch", [1, 2])
def test_resample(cuda, rates, ch):
    g, o = _pair(_rand(20011 * ch, 12), sr=rates[0], ch=ch)
    _same(g.resample(rates[1]), o.resample(rates[1]), "resample %s" % (rates,))
    # small and extreme values: negative sums a hair below an integer, where C's truncation
    # toward zero followed by >>16 differs from a plain floor
    rng = np.random.default_rng(5)
    for data in (rng.integers(-3, 4, 20011 * ch), rng.choice([-32768, 32767, -1, 0, 1], 20011 * ch)):
        g, o = _pair(data.astype(np.int16), sr=rates[0], ch=ch)
        _same(g.resample(rates[1]), o.resample(rates[1]), "resample %s small/extreme values" % (rates,))


@pytest.mark.parametrize("in_off,out_off", [(0, 0), (1, 0), (0, 1), (1, 1), (3, 2)])
def test_ratecv_views_at_odd_offsets(cuda, in_off, out_off):
    """Stereo frames that start on a 2-byte (not 4-byte) boundary take the scalar load/store
    path; lengths that end inside a warp's chunk take the guarded path.  Straight through the C ABI."""
    import torch
    from oracle import coracle as co
    from synthesizer_b200 import _cabi as C, _device as D
    for nin in (1, 2, 31, 257, 8191, 8192 * 3 + 5):
        src = _rand(2 * nin + 8, 50 + nin)
        dsrc = torch.from_numpy(src).cuda()
        want = co.pcm16_ratecv(src[in_off:in_off + 2 * nin], 2, 44100, 48000)
        nout = C.lib.sb_pcm16_ratecv_frames(nin, 44100, 48000)
        assert want.size == 2 * nout
        out = torch.full((2 * nout + 8,), 12345, dtype=torch.int16, device="cuda")
        C.call("sb_pcm16_ratecv", D.ptr(dsrc[in_off:]), nin, 2, 44100, 48000, D.ptr(out[out_off:]), D.stream())
        got = out.cpu().numpy()
        assert np.array_equal(got[out_off:out_off + 2 * nout], want), (nin, in_off, out_off)
        assert (got[:out_off] == 12345).all() and (got[out_off + 2 * nout:] == 12345).all(), "wrote outside"


def test_speed(cuda):
    g, o = _pair(_rand(30000, 13))
    _same(g.speed(1.37), o.speed(1.37), "speed")
    _same(g.speed(0.5), o.speed(0.5), "speed")


def test_fades(cuda):
    g, o = _pair(_rand(44100, 14))
    _same(g.fadeout(0.37, 0.1), o.fadeout(0.37, 0.1), "fadeout")
    _same(g.fadein(0.21, 0.05), o.fadein(0.21, 0.05), "fadein")
    _same(g.fadeout(5.0), o.fadeout(5.0), "fadeout longer than sample")
    g, o = _pair(_rand(44100 * 2, 15), ch=2)
    _same(g.fadein(0.5), o.fadein(0.5), "fadein stereo")
    _same(g.fadeout(0.25), o.fadeout(0.25), "fadeout stereo")


def test_envelope(cuda):
    g, o = _pair(_rand(44100, 16))
    _same(g.envelope(0.1, 0.2, 0.6, 0.3), o.envelope(0.1, 0.2, 0.6, 0.3), "envelope")
    g, o = _pair(_rand(22050, 17))
    _same(g.envelope(0.0, 0.1, 1.0, 0.0), o.envelope(0.0, 0.1, 1.0, 0.0), "envelope degenerate")


def test_echo(cuda):
    g, o = _pair(_rand(44100, 18, full=False))
    _same(g.echo(0.3, 4, 0.11, 0.6), o.echo(0.3, 4, 0.11, 0.6), "echo")


def test_huge_factors_and_ragged_lengths(cuda):
    """Factors above 16384 take the kernels' general (converting) path, small ones the
    conversion-free one; frame counts that are not a multiple of 4 exercise the scalar tails of
    the vectorised mono / stereo kernels.  Both must equal audioop."""
    for nfr in (1, 2, 3, 5, 4099):
        g, o = _pair(_rand(2 * nfr, 40 + nfr), ch=2)
        for lf, rf in ((0.5, 0.5), (1.0, -1.0), (16384.0, 1.0), (30000.0, -29999.5), (-1e6, 1e-6)):
            _same(g.copy().mono(lf, rf), o.copy().mono(lf, rf), "mono %r %r n=%d" % (lf, rf, nfr))
        m, mo = _pair(_rand(nfr, 41 + nfr))
        for lf, rf in ((1.0, 1.0), (0.7, 1.4), (-16384.0, 16384.0), (1e5, -1e5)):
            _same(m.copy().stereo(lf, rf), mo.copy().stereo(lf, rf), "stereo %r %r n=%d" % (lf, rf, nfr))
        for f in (16384.0, -16384.0, 16384.5, 70000.25, -1e9, 1e300):
            _same(m.copy().amplify(f), mo.copy().amplify(f), "amplify %r n=%d" % (f, nfr))
    g, o = _pair(_rand(30011, 42))
    for target in (0.0, 0.25, -0.5, 1.0):          # above 1 upstream's array('h') raises OverflowError
        _same(g.copy().fadeout(0.3, target), o.copy().fadeout(0.3, target), "fadeout to %r" % target)
        _same(g.copy().fadein(0.3, target), o.copy().fadein(0.3, target), "fadein from %r" % target)


def test_channels(cuda):
    g, o = _pair(_rand(20000, 19), ch=2)
    _same(g.copy().mono(), o.copy().mono(), "mono")
    _same(g.copy().mono(0.3, 0.9), o.copy().mono(0.3, 0.9), "mono factors")
    _same(g.copy().left(), o.copy().left(), "left")
    _same(g.copy().right(), o.copy().right(), "right")
    _same(g.copy().stereo(0.5, 1.2), o.copy().stereo(0.5, 1.2), "stereo of stereo")
    m, mo = _pair(_rand(10000, 20))
    _same(m.copy().stereo(), mo.copy().stereo(), "stereo")
    _same(m.copy().stereo(0.7, 1.4), mo.copy().stereo(0.7, 1.4), "stereo factors")
    _same(m.copy().pan(-0.4), mo.copy().pan(-0.4), "pan")


def test_normalize_and_stereo_mix(cuda):
    g, o = _pair(_rand(22050, 40), sr=22050, ch=1)
    _same(g.normalize(), o.normalize(), "normalize mono 22050")
    g, o = _pair(_rand(9600 * 2, 41), sr=48000, ch=2)
    _same(g.normalize(), o.normalize(), "normalize stereo 48000")
    g, o = _pair(_rand(8000, 42, full=False), sr=8000)
    h, ho = _pair(_rand(3000, 43, full=False), sr=8000)
    _same(g.copy().stereo_mix(h, "L", 0.7, 0.1), o.copy().stereo_mix(ho, "L", 0.7, 0.1), "stereo_mix L")
    _same(g.copy().stereo().stereo_mix(h, "R", 1.3, 0.0), o.copy().stereo().stereo_mix(ho, "R", 1.3, 0.0), "stereo_mix R")


def test_widths(cuda):
    g, o = _pair(_rand(10000, 21))
    g.make_32bit()
    o.make_32bit()
    assert g.samplewidth == 4 and g.get_frames() == o._frames
    g.make_16bit(maximize_amplitude=False)
    o.make_16bit(maximize_amplitude=False)
    _same(g, o, "32 -> 16")


def test_ragged_lengths_and_unaligned_views_through_the_c_abi(cuda):
    """The 128-bit paths have scalar heads and tails: every length 0..40 at every int16 offset
    0..8 of a buffer, for the vectorised kernels, against the C oracle / stdlib audioop."""
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    from oracle import coracle as co
    rng = np.random.default_rng(77)
    base_a = rng.integers(-32768, 32768, 64).astype(np.int16)
    base_b = rng.integers(-32768, 32768, 64).astype(np.int16)
    base_f = (rng.random(64) * 2.2 - 1.1).astype(np.float32)
    da, db = torch.tensor(base_a, device="cuda"), torch.tensor(base_b, device="cuda")
    df = torch.tensor(base_f, device="cuda")
    for off in range(9):
        for n in list(range(0, 41)):
            a, b = da[off:off + n], db[off + 3:off + 3 + n]            # different alignments on purpose
            out = torch.full((64,), 12345, dtype=torch.int16, device="cuda")
            o = out[off + 1:off + 1 + n]
            C.call("sb_pcm16_add", D.ptr(a), D.ptr(b), n, D.ptr(o), D.stream())
            want = co.pcm16_add(base_a[off:off + n], base_b[off + 3:off + 3 + n]) if n else np.zeros(0, np.int16)
            got = out.cpu().numpy()
            assert np.array_equal(got[off + 1:off + 1 + n], want), ("add", off, n)
            assert (got[:off + 1] == 12345).all() and (got[off + 1 + n:] == 12345).all(), ("add wrote outside", off, n)
            out.fill_(12345)
            C.call("sb_pcm16_mul", D.ptr(a), n, 0.77, D.ptr(o), D.stream())
            want = co.pcm16_mul(base_a[off:off + n], 0.77) if n else np.zeros(0, np.int16)
            got = out.cpu().numpy()
            assert np.array_equal(got[off + 1:off + 1 + n], want), ("mul", off, n)
            assert (got[:off + 1] == 12345).all() and (got[off + 1 + n:] == 12345).all(), ("mul wrote outside", off, n)
            buf = da.clone()
            C.call("sb_pcm16_fadeout", D.ptr(buf[off:]), n, 0.25, D.stream())
            want = base_a.copy()
            if n:
                want[off:off + n] = co.pcm16_fadeout(base_a[off:off + n], 0.25)
            assert np.array_equal(buf.cpu().numpy(), want), ("fadeout", off, n)
            out.fill_(12345)
            ovf = torch.zeros(1, dtype=torch.int64, device="cuda")
            C.call("sb_f32_to_pcm16", D.ptr(df[off:]), n, 20000.0, D.ptr(o), D.ptr(ovf), D.stream())
            want = np.trunc(base_f[off:off + n].astype(np.float32) * np.float32(20000.0)).astype(np.int64)
            got = out.cpu().numpy()
            assert np.array_equal(got[off + 1:off + 1 + n], want), ("to_pcm16", off, n)
            assert (got[:off + 1] == 12345).all() and (got[off + 1 + n:] == 12345).all(), ("to_pcm16 wrote outside", off, n)
            assert int(ovf.item()) == 0


def test_empty_samples_everywhere(cuda):
    import synthesizer_b200 as sb
    e = sb.Sample.from_raw_frames(b"", 2, 44100, 1)
    for name, op in [("amplify", lambda s: s.amplify(0.5)), ("fadein", lambda s: s.fadein(1.0)),
                     ("fadeout", lambda s: s.fadeout(1.0)), ("reverse", lambda s: s.reverse()),
                     ("bias", lambda s: s.bias(5)), ("resample", lambda s: s.resample(48000)),
                     ("envelope", lambda s: s.envelope(0.1, 0.1, 0.5, 0.1)), ("echo", lambda s: s.echo(1.0, 3, 0.1, 0.5)),
                     ("stereo", lambda s: s.stereo()), ("amplify_max", lambda s: s.amplify_max()),
                     ("make_32bit", lambda s: s.make_32bit()), ("clip", lambda s: s.clip(0, 0)),
                     ("add_silence", lambda s: s.add_silence(0.01)), ("delay", lambda s: s.delay(0.01))]:
        g, o = _pair([])
        assert op(g) is g
        op(o)
        assert (g.samplerate, g.nchannels, g.samplewidth) == (o.samplerate, o.nchannels, o.samplewidth), name
        assert g.get_frames() == o._frames, name
    assert e.copy().mix_at_many([]).get_frames() == b""
    x = sb.Sample.from_raw_frames(np.arange(10, dtype="<i2").tobytes(), 2, 44100, 1)
    assert e.copy().mix(x).get_frames() == x.get_frames()
    assert x.copy().mix(e).get_frames() == x.get_frames()
    assert e.copy().mix_at(0.0001, x).get_frames()[-20:] == x.get_frames()
    import synthesizer_b200.oscillators as so_
    assert so_.Sine(440, samplerate=44100).render(0).numel() == 0


def test_32bit_samples(cuda):
    """Sample width 4: every arithmetic routine against stdlib audioop with size 4."""
    import synthesizer_b200 as sb
    from oracle.sample_oracle import Sample as OS
    rng = np.random.default_rng(31)

    def pair(n, ch=1, sr=44100, full=True):
        lim = 2 ** 31 if full else 2 ** 29
        x = rng.integers(-lim, lim, size=n * ch, dtype=np.int64)
        x[:4] = [2 ** 31 - 1, -2 ** 31, -2 ** 31 + 1, 0]
        raw = x.astype("<i4").tobytes()
        return sb.Sample.from_raw_frames(raw, 4, sr, ch), OS.from_raw_frames(raw, 4, sr, ch)

    def same(g, o, what):
        assert (g.samplerate, g.nchannels, g.samplewidth) == (o.samplerate, o.nchannels, o.samplewidth), what
        dt = "<i4" if g.samplewidth == 4 else "<i2"
        a, b = np.frombuffer(g.get_frames(), dt), np.frombuffer(o._frames, dt)
        assert a.shape == b.shape, what
        assert np.array_equal(a, b), "%s: %d of %d differ" % (what, int((a != b).sum()), a.size)

    g, o = pair(20001)
    assert g.maximum == o.maximum
    for f in (0.5, 1.7, -1.0, 0.333, 1e-3):
        same(g.copy().amplify(f), o.copy().amplify(f), "amplify %r" % f)
    g2, o2 = pair(20001)
    same(g.copy().mix(g2), o.copy().mix(o2), "mix")
    same(g.copy().mix_at(0.1, g2), o.copy().mix_at(0.1, o2), "mix_at")
    same(g.copy().mix_at_many([(0.05, g2), (0.2, g2)]), o.copy().mix_at(0.05, o2).mix_at(0.2, o2), "mix_at_many")
    same(g.copy().bias(123456789), o.copy().bias(123456789), "bias")
    same(g.copy().reverse(), o.copy().reverse(), "reverse")
    same(g.copy().fadeout(0.2, 0.1), o.copy().fadeout(0.2, 0.1), "fadeout")
    same(g.copy().fadein(0.2, 0.1), o.copy().fadein(0.2, 0.1), "fadein")
    same(g.copy().resample(48000), o.copy().resample(48000), "resample")
    same(g.copy().envelope(0.05, 0.1, 0.6, 0.1), o.copy().envelope(0.05, 0.1, 0.6, 0.1), "envelope")
    s, so_ = pair(10000, ch=2, full=False)
    same(s.copy().mono(0.4, 0.7), so_.copy().mono(0.4, 0.7), "mono")
    same(s.copy().resample(22050), so_.copy().resample(22050), "resample stereo")
    same(g.copy().stereo(0.7, 1.2), o.copy().stereo(0.7, 1.2), "stereo")
    same(g.copy().amplify_max(), o.copy().amplify_max(), "amplify_max")
    same(g.copy().make_16bit(maximize_amplitude=False), o.copy().make_16bit(maximize_amplitude=False), "make_16bit")
    m, mo = _pair(_rand(5000, 32))
    same(m.copy().make_32bit(), mo.copy().make_32bit(), "make_32bit")
    same(m.copy().make_32bit(scale_amplitude=False), mo.copy().make_32bit(scale_amplitude=False), "make_32bit unscaled")
    ws = sb.WaveSynth(samplerate=44100, samplewidth=4).sine(440, 0.1)
    assert ws.samplewidth == 4 and len(ws) == 4410 and 2 ** 31 * 0.99 < ws.maximum < 2 ** 31


def test_structure_ops(cuda):
    g, o = _pair(_rand(44100, 22))
    _same(g.clip(0.1, 0.7), o.clip(0.1, 0.7), "clip")
    cg, co_ = g.split(0.25), o.split(0.25)
    _same(g, o, "split head")
    _same(cg, co_, "split tail")
    _same(g.add_silence(0.1), o.add_silence(0.1), "silence end")
    _same(g.add_silence(0.05, at_start=True), o.add_silence(0.05, at_start=True), "silence start")
    _same(g.join(cg), o.join(co_), "join")
    _same(g.delay(0.1), o.delay(0.1), "delay")
    _same(g.delay(-0.05), o.delay(-0.05), "delay negative")
    _same(g.delay(0.02, keep_length=True), o.delay(0.02, keep_length=True), "delay keep")
    _same(g.delay(-0.02, keep_length=True), o.delay(-0.02, keep_length=True), "delay negative keep")


def test_modulate_amp(cuda):
    import synthesizer_b200 as sb
    from oracle import synth_oracle as so
    g, o = _pair(_rand(20000, 23, full=False))
    g.modulate_amp(sb.Sine(5, 0.5, bias=0.5, samplerate=44100))
    o.modulate_amp(so.Sine(5, 0.5, bias=0.5, samplerate=44100))
    a = np.frombuffer(g.get_frames(), dtype="<i2").astype(int)
    b = np.frombuffer(o._frames, dtype="<i2").astype(int)
    assert np.abs(a - b).max() <= 1          # the modulator is float: +-1 LSB
    g, o = _pair(_rand(5000, 24, full=False))
    mod = [0, 50, 100, 50, 0, -50, -100, -50]
    _same(g.modulate_amp(mod), o.modulate_amp(mod), "modulate by list")
    # ratios float32 cannot represent: 70/100 * 10 is 7 in fp64 and 6 in float32
    frames = np.tile(np.array([10, -10, 30, 7, 32767, -32768, 1], dtype=np.int16), 300)
    g, o = _pair(frames)
    mod = [0, 30, 70, 100, -70]
    _same(g.modulate_amp(mod), o.modulate_amp(mod), "modulate by list, fp64 factors")
    g, o = _pair(frames)
    import synthesizer_b200 as sb2
    from oracle import sample_oracle as sa
    ms = np.array([3, 7, -11, 13], dtype="<i2").tobytes()
    _same(g.modulate_amp(sb2.Sample.from_raw_frames(ms, 2, 44100, 1)),
          o.modulate_amp(sa.Sample.from_raw_frames(ms, 2, 44100, 1)), "modulate by Sample")


def test_wav_roundtrip(cuda, tmp_path):
    import synthesizer_b200 as sb
    g, o = _pair(_rand(12345 * 2, 25), sr=48000, ch=2)
    p = str(tmp_path / "x.wav")
    g.write_wav(p)
    back = sb.Sample(p)
    assert (back.samplerate, back.nchannels, back.samplewidth) == (48000, 2, 2)
    assert back.get_frames() == o._frames


def test_locked_sample_refuses(cuda):
    g, _ = _pair(_rand(100, 26))
    g.lock()
    with pytest.raises(AssertionError):
        g.amplify(0.5)


def test_mix_at_many_sharded_slices_equal_the_unsharded_result(cuda):
    """Destination-range sharding (config 4 across GPUs, no collective): the slice a rank mixes is
    bit-identical to the same range of the whole mix.  The ranks are played one after another
    on this GPU (shard=(rank, world))."""
    import synthesizer_b200 as sb
    rng = np.random.default_rng(12)
    sr, total = 48000, 48000 * 3
    base = rng.integers(-9000, 9000, total).astype(np.int16)
    hits = [(float(np.float32(rng.uniform(0, 2.9))),
             sb.Sample.from_raw_frames(rng.integers(-30000, 30000, int(rng.integers(1, 9000))).astype("<i2").tobytes(), 2, sr, 1))
            for _ in range(300)]
    hits.append((2.99, sb.Sample.from_raw_frames(rng.integers(-30000, 30000, 5000).astype("<i2").tobytes(), 2, sr, 1)))  # grows it
    whole = sb.Sample.from_raw_frames(base.tobytes(), 2, sr, 1).mix_at_many(hits)
    want = np.frombuffer(whole.get_frames(), dtype="<i2")
    for world in (2, 3, 8):
        got = []
        for rank in range(world):
            part = sb.Sample.from_raw_frames(base.tobytes(), 2, sr, 1).mix_at_many(hits, shard=(rank, world))
            lo, hi, tot = part.shard_range
            assert tot == want.size and (lo, hi) == sb.Sample.shard_bounds(tot, rank, world)
            got.append(np.frombuffer(part.get_frames(), dtype="<i2"))
            assert got[-1].size == hi - lo
        assert np.array_equal(np.concatenate(got), want), "world %d" % world


def test_mix_at_many_bad_input_stays_inside_its_buffers(cuda):
    """Raw C-ABI callers: a negative destination offset is left out (and counted), a scratch too
    small for the per-tile lists is refused per tile (flagged) - neither writes out of bounds."""
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    rng = np.random.default_rng(5)
    total, nsrc = 40000, 64
    dst = torch.zeros(total, dtype=torch.int16, device="cuda")
    lens = rng.integers(100, 3000, nsrc).astype(np.int64)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    starts = rng.integers(0, total - 3000, nsrc).astype(np.int64)
    starts[5] = -17                                                # bad
    pool = torch.from_numpy(rng.integers(-3000, 3000, int(lens.sum())).astype(np.int16)).cuda()
    idx = torch.tensor(np.stack([offs, lens, starts]), dtype=torch.int64, device="cuda")
    nscr = C.lib.sb_pcm16_mix_at_many_scratch_bytes(total, nsrc, int(lens.sum()))
    guard = 4096
    scr = torch.full((nscr + guard,), 0x5A, dtype=torch.uint8, device="cuda")
    C.call("sb_pcm16_mix_at_many", D.ptr(dst), total, D.ptr(pool), D.ptr(idx[0]), D.ptr(idx[1]), D.ptr(idx[2]), nsrc,
           D.ptr(scr), nscr, D.stream())
    status = scr[:16].view(torch.int32).cpu().numpy()
    assert status[1] == 1 and status[2] == 0
    assert bool((scr[nscr:] == 0x5A).all())
    # the same mix without the bad source is what came out
    want = np.zeros(total, dtype=np.int64)
    p = pool.cpu().numpy().astype(np.int64)
    for s in range(nsrc):
        if starts[s] >= 0:
            want[starts[s]:starts[s] + lens[s]] = np.clip(want[starts[s]:starts[s] + lens[s]] + p[offs[s]:offs[s] + lens[s]], -32768, 32767)
    assert np.array_equal(dst.cpu().numpy().astype(np.int64), want)
    # a scratch with room for the fixed part and three list entries only
    ntiles = (total + 2047) // 2048
    small = ((3 * ntiles + 4) * 4 + 15) // 16 * 16 + 3 * 32       # (a list entry is 32 bytes)
    scr2 = torch.full((small + guard,), 0x5A, dtype=torch.uint8, device="cuda")
    dst.zero_()
    C.call("sb_pcm16_mix_at_many", D.ptr(dst), total, D.ptr(pool), D.ptr(idx[0]), D.ptr(idx[1]), D.ptr(idx[2]), nsrc,
           D.ptr(scr2), small, D.stream())
    assert scr2[:16].view(torch.int32).cpu().numpy()[2] == 1
    assert bool((scr2[small:] == 0x5A).all())


def test_mix_at_many_with_a_placement_table(cuda):
    """The reusable table gives the same bytes as the list of placements, on every use."""
    import synthesizer_b200 as sb
    rng = np.random.default_rng(21)
    sr = 44100
    base = rng.integers(-9000, 9000, sr).astype(np.int16)
    hits = [(float(np.float32(rng.uniform(0, 1.2))),
             sb.Sample.from_raw_frames(rng.integers(-30000, 30000, int(rng.integers(0, 5000))).astype("<i2").tobytes(), 2, sr, 1))
            for _ in range(100)]
    want = np.frombuffer(sb.Sample.from_raw_frames(base.tobytes(), 2, sr, 1).mix_at_many(hits).get_frames(), dtype="<i2")
    table = sb.Sample.from_raw_frames(b"", 2, sr, 1).placement_table(hits)
    for _ in range(2):
        got = sb.Sample.from_raw_frames(base.tobytes(), 2, sr, 1).mix_at_many(table)
        assert np.array_equal(np.frombuffer(got.get_frames(), dtype="<i2"), want)
    with pytest.raises(ValueError):
        sb.Sample.from_raw_frames(base.tobytes(), 2, sr, 2).mix_at_many(table)


def _pair24(values, sr=44100, ch=1):
    import synthesizer_b200 as sb
    from oracle.sample_oracle import Sample as OS
    v = np.asarray(values, dtype=np.int64) & 0xFFFFFF
    raw = np.stack([v & 0xFF, (v >> 8) & 0xFF, (v >> 16) & 0xFF], axis=1).astype(np.uint8).tobytes()
    return sb.Sample.from_raw_frames(raw, 3, sr, ch), OS.from_raw_frames(raw, 3, sr, ch)


def _same24(g, o, what):
    assert (g.samplerate, g.nchannels, g.samplewidth) == (o.samplerate, o.nchannels, o.samplewidth), what
    assert g.get_frames() == o._frames, what


def test_24_bit_frames_follow_audioop_size_3(cuda, tmp_path):
    """3-byte frames: every audioop-backed operation, bit for bit against the real module."""
    import synthesizer_b200 as sb
    rng = np.random.default_rng(33)
    full = rng.integers(-(1 << 23), 1 << 23, 6000)
    full[:4] = [(1 << 23) - 1, -(1 << 23), -(1 << 23) + 1, 0]
    g, o = _pair24(full)
    _same24(g, o, "round trip")
    assert len(g) == len(o) and g.maximum == o.maximum
    for factor in (0.5, 1.7, -1.0, 0.333, 2.75):
        g, o = _pair24(full)
        _same24(g.amplify(factor), o.amplify(factor), "amplify %g" % factor)
    g, o = _pair24(full)
    g2, o2 = _pair24(np.roll(full, 7))
    _same24(g.mix(g2), o.mix(o2), "mix (saturates at 24 bits)")
    g, o = _pair24(full[:5000])
    _same24(g.mix_at(0.01, g2), o.mix_at(0.01, o2), "mix_at")
    g, o = _pair24(full)
    _same24(g.bias(3_000_000), o.bias(3_000_000), "bias (wraps at 24 bits)")
    g, o = _pair24(full, ch=2)
    _same24(g.mono(0.8, 0.7), o.mono(0.8, 0.7), "mono")
    g, o = _pair24(full)
    _same24(g.stereo(1.3, 0.6), o.stereo(1.3, 0.6), "stereo")
    g, o = _pair24(full, ch=2)
    _same24(g.reverse(), o.reverse(), "reverse")
    for rate in (48000, 22050, 44101):
        g, o = _pair24(full, ch=2)
        _same24(g.resample(rate), o.resample(rate), "resample %d" % rate)
    g, o = _pair24(full)
    _same24(g.amplify_max(), o.amplify_max(), "amplify_max")
    g, o = _pair24(full)
    g.make_32bit(); o.make_32bit()
    assert g.samplewidth == 4 and g.get_frames() == o._frames
    g, o = _pair24(full // 3)
    g.make_16bit(); o.make_16bit()
    assert g.samplewidth == 2 and g.get_frames() == o._frames
    # WAV round trip, and what upstream cannot do at this width either
    g, o = _pair24(full, sr=48000, ch=2)
    path = str(tmp_path / "x24.wav")
    g.write_wav(path)
    back = sb.Sample(path)
    assert (back.samplewidth, back.nchannels, back.samplerate) == (3, 2, 48000) and back.get_frames() == o._frames
    with pytest.raises(ValueError):
        g.fadeout(0.01)
    with pytest.raises(ValueError):
        g.get_frame_array()


def test_pan_with_an_lfo(cuda):
    import synthesizer_b200 as sb
    from oracle import synth_oracle as so
    for ch in (1, 2):
        g, o = _pair(np.random.default_rng(41).integers(-15000, 15000, 4000 * ch), sr=48000, ch=ch)
        g.pan(lfo=sb.Sine(3.0, 0.8, samplerate=48000))
        o.pan(lfo=so.Sine(3.0, 0.8, samplerate=48000))
        assert g.nchannels == 2 and len(g) == 4000
        a = np.frombuffer(g.get_frames(), dtype="<i2").astype(int)
        b = np.frombuffer(o._frames, dtype="<i2").astype(int)
        assert np.abs(a - b).max() <= 1          # the panning values are float32 on the device
