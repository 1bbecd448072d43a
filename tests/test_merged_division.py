"""The arithmetic behind fheb200_eval_mult_rescaled (Context::relinearize_reduced): ModDown by R = P * q_last with an exact
(Halevi-Polyakov-Shoup) lift, restated with Python integers.  For X given by its residues modulo the ordinary limbs {q_i} and the
special limbs {r_j} (R = prod r_j):

    y_j   = [X * (R / r_j)^-1] mod r_j                          (what the inverse NTT's custom scaling leaves in limb j)
    u     = round(sum_j y_j / r_j)   in float64                 (csrc/baseconv.cu: bc_lift_overflow)
    lift  = sum_j y_j * (R / r_j) - u * R  ==  the centred representative of X mod R   (the conversion table's extra source, -R mod q_i)
    out_i = (X_i - lift) * R^-1 mod q_i  ==  round(X / R) mod q_i                       (the fused ModDown finish)

A plain fast base conversion (no u) is off by up to len(r) multiples of R: harmless when a rescale follows, not when this quotient
is the result - which is why the merged division needs the exact form (DESIGN.md, round-2 results)."""
import numpy as np

Q = [1152921504606584833, 1099512938497, 1099510054913, 1099507695617]          # ordinary limbs (60-bit first, 40-bit scaling)
R_LIMBS = [1099506515969, 1152921504606748673, 1152921504606683137]          # q_last followed by the special limbs ("P")


def test_exact_lift_gives_the_rounded_quotient():
    from math import gcd
    r, q = list(R_LIMBS), list(Q)
    assert all(gcd(a, b) == 1 for i, a in enumerate(r + q) for b in (r + q)[i + 1:])   # CRT only needs pairwise coprime moduli
    R = 1
    for m in r:
        R *= m
    Qp = 1
    for m in q:
        Qp *= m
    rng = np.random.default_rng(3)
    worst_plain = 0
    for _ in range(400):
        # a "ciphertext coefficient": any integer in (-Q'R/2, Q'R/2), here with a wide spread of magnitudes
        X = int(rng.integers(-2**62, 2**62)) * int(rng.integers(1, 2**62)) * int(rng.integers(1, 2**40)) % (Qp * R) - (Qp * R) // 2
        y = [(X % m) * pow(R // m, -1, m) % m for m in r]
        u = int(np.rint(sum(np.float64(yj) / np.float64(m) for yj, m in zip(y, r))))
        lift_fast = sum(yj * (R // m) for yj, m in zip(y, r))
        lift = lift_fast - u * R
        centred = X % R
        if centred > R // 2:
            centred -= R
        assert lift == centred
        want = (X - centred) // R                        # round(X / R), exact
        assert abs(want - X / R) <= 0.5 + 1e-9
        for m in q:
            got = ((X % m) - (lift % m)) * pow(R, -1, m) % m
            assert got == want % m
        worst_plain = max(worst_plain, abs((X - lift_fast) // R - want))
    assert 1 <= worst_plain <= len(r)                    # the uncorrected conversion undershoots by up to len(r) units
