"""Exactness of the FP64 modular multiplication of the NTT's double-precision path (csrc/ntt.cu fp_mulmod), emulated on
the CPU: IEEE double products by Python floats, fused multiply-adds by exact integer arithmetic rounded once."""
import random
from fractions import Fraction

MAGIC = 6755399441055744.0  # 1.5 * 2^52


def rn(x):
    return float(Fraction(x))


def fp_mulmod(y, w, qinv, q):
    h = float(y) * float(w)                                       # DMUL, rounded
    c = rn(Fraction(h) * Fraction(qinv) + Fraction(MAGIC)) - MAGIC  # rint(h/q): DFMA with the magic constant + DADD
    lo = Fraction(int(y)) * int(w) - Fraction(int(h))             # fma(y, w, -h): exact error term
    assert rn(lo) == lo
    d = Fraction(int(h)) - Fraction(int(c)) * q                   # fma(-c, q, h): small integer, exact
    assert rn(d) == d
    g = rn(d) + rn(lo)
    assert g == int(d) + int(lo)
    return g


def test_fp_mulmod_is_exact_and_centred():
    rnd = random.Random(1)
    worst = 0.0
    for bits in (36, 40, 41, 42):
        for _ in range(20000):
            q = rnd.getrandbits(bits) | (1 << (bits - 1)) | 1
            w = rnd.randrange(q)
            qinv = rn(Fraction(1, q))
            bound = rnd.choice([1, 4, 20, 256, 512])
            y = float(rnd.randrange(-bound * q, bound * q))
            if abs(y) >= 2.0 ** 50:
                continue
            g = fp_mulmod(y, w, qinv, q)
            assert (int(g) - int(y) * w) % q == 0
            worst = max(worst, abs(g) / q)
    assert worst < 0.75


def test_u64_double_round_trip_bit_tricks():
    two52 = 4503599627370496.0
    for x in (0, 1, 2 ** 41 - 1, 2 ** 42 - 1, 2 ** 52 - 1):
        assert (float(x) + two52) - two52 == float(x)
