"""PcmStreamer: render-ahead chunked PCM equals the one-shot render, chunk by chunk."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _chain(sb, sr):
    lfo = sb.Sine(5.0, 0.03, samplerate=sr)
    car = sb.Sawtooth(220.0, 0.45, fm_lfo=lfo, samplerate=sr)
    return sb.EchoFilter(sb.EnvelopeFilter(car, 0.05, 0.1, 0.5, 0.6, 0.3), 0.1, 3, 0.07, 0.5)


def test_streamer_matches_one_shot(cuda):
    import synthesizer_b200 as sb
    from synthesizer_b200.streaming import PcmStreamer
    sr, total, chunk = 48000, 100_000, 16384
    whole = np.frombuffer(sb.Sample.from_oscillator(_chain(sb, sr), total / sr).get_frames(), dtype="<i2")
    chunks = list(PcmStreamer(_chain(sb, sr), chunk_frames=chunk, max_frames=total, ahead=2))
    assert [len(c) // 2 for c in chunks] == [chunk] * (total // chunk) + [total % chunk]
    got = np.frombuffer(b"".join(chunks), dtype="<i2")
    assert got.shape == whole.shape
    assert np.abs(got.astype(int) - whole.astype(int)).max() <= 1


def test_streamer_ends_with_a_finite_oscillator_and_reports_overflow(cuda):
    import synthesizer_b200 as sb
    from synthesizer_b200.streaming import PcmStreamer
    sr = 44100
    osc = sb.EnvelopeFilter(sb.Sine(440, 0.5, samplerate=sr), 0.01, 0.02, 0.05, 0.5, 0.02, stop_at_end=True)
    n = sum(len(c) // 2 for c in PcmStreamer(osc, chunk_frames=1000))
    assert n == osc._length() and 4400 <= n <= 4420
    with pytest.raises(OverflowError):
        list(PcmStreamer(sb.Sine(440, 1.5, samplerate=sr), chunk_frames=1000, max_frames=3000))
