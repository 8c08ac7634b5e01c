"""
Checks, in exact rational arithmetic, that the fp64 quotient the fade kernels compute with a
correctly rounded reciprocal and two fused correction steps equals the IEEE quotient CPython
computes for Sample.fadeout/fadein's  k*decrease/numsamples  (no GPU needed):

    y  = RN(1/b)          q0 = RN(a*y)
    r0 = RN(a - b*q0)     q1 = RN(q0 + r0*y)        (fused: one rounding each)
    r1 = RN(a - b*q1)     q2 = RN(q1 + r1*y)

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
    r1 = fma(-b, q1, a)
    return fma(r1, y, q1), q1


def main():
    rnd = random.Random(7)
    first_step_misses = 0
    cases = 0
    for trial in range(400):
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
            got, q1 = div_by_reciprocal(a, b)
            cases += 1
            first_step_misses += q1 != want
            assert got == want, (k, dec, n, got, want)
    print("ok: %d quotients identical (%d needed the second correction step)" % (cases, first_step_misses))


if __name__ == "__main__":
    main()
