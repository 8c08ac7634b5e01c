"""
Checks, in exact rational arithmetic, that the fp64 quotient the fade kernels compute with a
correctly rounded reciprocal and ONE fused correction step equals the IEEE quotient CPython
computes for Sample.fadeout/fadein's  k*decrease/numsamples  (no GPU needed):

    y  = RN(1/b)          q0 = RN(a*y)
    r0 = RN(a - b*q0)     q1 = RN(q0 + r0*y)        (fused: one rounding each)

(Markstein's division theorem: with y the correctly rounded reciprocal and q0 within an ulp of
a/b, q1 is the correctly rounded quotient unless b's significand is all ones, which a sample
count below 2^53 - 1 never is.  The kernels did a second step until round 2; it never changed a
quotient, here or on the GPU.)

Run: python tools/fma_division_check.py
"""
import random
from fractions import Fraction


def rn(x):
    """Correctly rounded double of an exact rational."""
    return x.numerator / x.denominator


def fma(a, b, c):
    return rn(Fraction(a) * Fraction(b) + Fraction(c))


def div_by_reciprocal(a, b):
    y = 1.0 / b
    q0 = a * y
    r0 = fma(-b, q0, a)
    q1 = fma(r0, y, q0)
    return q1, q0


def main():
    rnd = random.Random(7)
    first_step_misses = 0
    cases = 0
    for trial in range(1500):
        n = rnd.choice([1, 2, 3, 7, 4800, 48000, 44100, 96000, 2 ** 20 - 1, 2 ** 31 - 1, 28_800_000,
                        rnd.randrange(1, 1 << 36), rnd.randrange(1, 1 << 20)])
        dec = rnd.choice([1.0, 0.5, 1.0 - 0.1, 1.0 - 1 / 3, rnd.random(), rnd.random() * 1e-3, 1.0 - 1e-9])
        b = float(n)
        ks = {0, 1, n - 1, n // 2, n // 3} | {rnd.randrange(0, n) for _ in range(300)}
        for k in ks:
            if k < 0:
                continue
            a = float(k) * dec
            want = a / b
            got, q0 = div_by_reciprocal(a, b)
            cases += 1
            first_step_misses += q0 != want
            assert got == want, (k, dec, n, got, want)
    print("ok: %d quotients identical (%d of them differ before the correction step)" % (cases, first_step_misses))


if __name__ == "__main__":
    main()
