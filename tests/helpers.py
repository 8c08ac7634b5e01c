"""Shared helpers of the parity tests (the oracle is only ever the checker)."""
import numpy as np


def pcm16_of(values, scale=32767.0):
    """int(v*scale) with truncation toward zero, from fp64 oracle values."""
    x = np.asarray(values, dtype=np.float64) * scale
    return np.trunc(x).astype(np.int64)


def assert_within_lsb(got, want, tol=1, what=""):
    got = np.asarray(got).astype(np.int64)
    want = np.asarray(want).astype(np.int64)
    assert got.shape == want.shape, "%s: shape %s vs %s" % (what, got.shape, want.shape)
    diff = np.abs(got - want)
    worst = int(diff.max()) if diff.size else 0
    if worst > tol:
        idx = int(diff.argmax())
        raise AssertionError("%s: %d samples differ by more than %d LSB (worst %d at index %d: got %d want %d)"
                             % (what, int((diff > tol).sum()), tol, worst, idx, got[idx], want[idx]))


def dev_pcm16(t, scale=32767.0):
    """Truncating float->int16 conversion of a CUDA float tensor through the C ABI."""
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    out = D.empty(t.numel(), torch.int16)
    ovf = D.zeros(1, torch.int64)
    C.call("sb_f32_to_pcm16", D.ptr(t), t.numel(), float(scale), D.ptr(out), D.ptr(ovf), D.stream())
    assert int(ovf.item()) == 0, "overflow in float->pcm16"
    return out.cpu().numpy()
