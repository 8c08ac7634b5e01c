"""
Checks the all-32-bit-integer form of audioop.ratecv's per-sample arithmetic (the one
pcm16_ratecv_kernel uses) against CPython's own audioop, on this interpreter.

  out = ((prev<<16)*d + (cur<<16)*(outrate-d)) / outrate   (fp64, truncated, then >>16)

reduces, for reduced rates below 2^21, to  cur + floor((prev-cur)*d / outrate)  plus a +1
fix-up when the sum is negative and its remainder is within outrate/65536 of a multiple.
Run: python tools/ratecv_formula_check.py
"""
import math
import sys
import warnings

import numpy as np

warnings.simplefilter("ignore")
import audioop  # noqa: E402


def fast(a, nch, inr, outr):
    g = math.gcd(inr, outr)
    inr //= g
    outr //= g
    a = a.reshape(-1, nch).astype(np.int64)
    nin = len(a)
    nout = ((nin - 1) * outr) // inr + 1
    j = np.arange(nout, dtype=np.int64)
    c = -((-j * inr) // outr)
    d = c * outr - j * inr
    cur = a[c]
    prev = np.where((c > 0)[:, None], a[np.maximum(c - 1, 0)], 0)
    diff = (prev - cur).astype(np.int32)
    rcp = np.float32(1.0) / np.float32(outr)
    w = (d.astype(np.float32) * rcp).astype(np.float32)
    # fp32 estimate biased down by more than its error: floor or floor-1 (the product is rounded
    # once, as the kernel's fused multiply-add does: float64 holds it exactly before the cast)
    x = (diff.astype(np.float64) * w[:, None].astype(np.float64) - 0.015625).astype(np.float32)
    est = np.floor(x).astype(np.int64)
    elo = diff.astype(np.int64) * d[:, None]
    r = (elo - est * outr) % 2 ** 32                                 # what 32-bit wrapping arithmetic keeps
    assert (r < 2 * outr).all()
    hi = r >= outr
    est = est + hi
    r = r - hi * outr
    qf = cur + est
    lim = outr - (-(-outr // 65536))
    fix = (qf < 0) & (r > lim)
    return (qf + fix).astype(np.int16).reshape(-1)


def fast_magic(a, nch, inr, outr):
    """The kernel's form for reduced output rates <= 16384: the same closed-form (c, d), and
    cur + floor((prev-cur)*d/outr) by one widening multiply and a shift (invariant division)."""
    g = math.gcd(inr, outr)
    inr //= g
    outr //= g
    assert outr <= 16384
    a = a.reshape(-1, nch).astype(np.int64)
    nout = ((len(a) - 1) * outr) // inr + 1
    j = np.arange(nout, dtype=np.int64)
    c = -((-j * inr) // outr)
    d = c * outr - j * inr
    cur = a[c]
    prev = np.where((c > 0)[:, None], a[np.maximum(c - 1, 0)], 0)
    L = 0
    while (1 << L) < outr:
        L += 1
    msh = 31 + L
    mg = ((1 << msh) + outr - 1) // outr
    assert mg < 2 ** 32
    n = ((prev - cur) * d[:, None] + 65536 * outr)
    assert (n >= 0).all() and (n < 2 ** 31).all()
    q = (n.astype(object) * mg) >> msh                       # 64-bit product in the kernel
    return (cur + q.astype(np.int64) - 65536).astype(np.int16).reshape(-1)


def main():
    rng = np.random.default_rng(1)
    pairs = [(44100, 48000), (48000, 44100), (22050, 48000), (48000, 8000), (44100, 44101), (8000, 96000),
             (11025, 44100), (1, 3), (3, 1), (2000000, 1999999), (1999999, 2000000), (7, 2097151), (2097151, 7),
             (96000, 44100), (65537, 65536), (131071, 131072), (1048573, 1048576), (32000, 48000), (48000, 16383),
             (16385, 16384), (12345, 16381), (44100, 22050), (5, 16384)]
    for inr, outr in pairs:
        for nch in (1, 2):
            for kind in range(3):
                n = max(2, min(20011, 400000 * inr // outr)) * nch      # bound the output length
                if kind == 0:
                    a = rng.integers(-32768, 32768, n).astype(np.int16)
                elif kind == 1:
                    a = rng.choice(np.array([-32768, 32767, -1, 0, 1], dtype=np.int16), n)
                else:
                    a = rng.integers(-3, 4, n).astype(np.int16)
                want = np.frombuffer(audioop.ratecv(a.tobytes(), 2, nch, inr, outr, None)[0], dtype=np.int16)
                got = fast(a, nch, inr, outr)
                assert len(want) == len(got), (inr, outr, len(want), len(got))
                bad = np.flatnonzero(want != got)
                assert bad.size == 0, (inr, outr, nch, kind, bad[:5], want[bad[:5]], got[bad[:5]])
                if outr // math.gcd(inr, outr) <= 16384:
                    got = fast_magic(a, nch, inr, outr)
                    bad = np.flatnonzero(want != got)
                    assert bad.size == 0, ("magic", inr, outr, nch, kind, bad[:5], want[bad[:5]], got[bad[:5]])
    print("ok: %d rate pairs" % len(pairs))


if __name__ == "__main__":
    sys.exit(main())
