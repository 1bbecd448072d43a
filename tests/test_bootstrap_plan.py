"""Host-side checks of the bootstrapping plan (fhe_sorting_b200/csrc/bootstrap_plan.cpp) in the clear: the grouped
CoeffsToSlots / SlotsToCoeffs diagonals against numpy's own special FFT, and the scaled-cosine EvalMod, on the exact
vectors a raised ciphertext would hold (message + q0 * I).  No GPU."""
import numpy as np
import pytest
from numpy.polynomial import chebyshev as Cheb

import fhe_sorting_b200.lib as F


def rot_group(s):
    r, p = [], 1
    for _ in range(s):
        r.append(p)
        p = p * 5 % (4 * s)
    return r


def special_matrix(s):
    xi = np.exp(2j * np.pi / (4 * s))
    return np.array([[xi ** ((g * k) % (4 * s)) for k in range(s)] for g in rot_group(s)])


def bitrev(s):
    lg = s.bit_length() - 1
    return np.array([int(format(i, "0%db" % lg)[::-1], 2) if lg else 0 for i in range(s)])


@pytest.mark.parametrize("s,budget,n", [(2, (1, 1), 1 << 10), (8, (2, 2), 1 << 12), (16, (4, 4), 1 << 10),
                                        (64, (3, 2), 1 << 12), (1024, (5, 5), 1 << 17)])
def test_cleartext_bootstrap(s, budget, n):
    rng = np.random.default_rng(s)
    q0, sf1, c = float((1 << 60) - 93), float((1 << 59) + 12345), 12
    info = F.bootstrap_plan_info(n, s, budget[0], budget[1], q0, sf1, c)
    K, r = info["K"], info["r"]
    assert info["depth"] == min(budget[0], s.bit_length() - 1) + 7 + r + min(budget[1], s.bit_length() - 1)
    g, T = n // (2 * s), q0 * 2.0 ** -c
    z = rng.uniform(0, 1, s) + 0j
    U = special_matrix(s)
    w = np.linalg.solve(U, z)
    I = rng.integers(-K + 1, K, 2 * s)
    t = np.concatenate([w.real, w.imag]) * T + q0 * I           # coefficients of the subring polynomial after ModRaise
    x0 = U @ ((t[:s] + 1j * t[s:]) * g / sf1)                     # slots after SubSum under the scale label sf1
    x = F.bootstrap_plan_apply(n, s, budget[0], budget[1], q0, sf1, 0, np.concatenate([x0, x0]), c)
    x = x + np.conj(x)
    br = bitrev(s)
    delta = 1.0 / (4 * K)
    want = np.concatenate([t[:s][br], t[s:][br]]) / (q0 * K)
    assert np.abs(x * (1 + delta) - want).max() < 1e-9
    # EvalMod as the engine runs it: u = x - shift, y = 2 u^2 - 1, even-part series in y, r double angles
    assert abs(info["shift"] - delta / (1 + delta)) < 1e-15 and len(info["cheb"]) == 45
    u = x.real - info["shift"]
    y = Cheb.chebval(2 * u * u - 1, np.concatenate([[info["cheb"][0] / 2], info["cheb"][1:]]))
    for _ in range(r):
        y = 2 * y * y - 1
    assert np.abs(y - np.sin(2 * np.pi * K * want.real)).max() < 1e-8
    v = F.bootstrap_plan_apply(n, s, budget[0], budget[1], q0, sf1, 1, y.astype(complex), c)
    v = v + np.roll(v, -s)
    # the sine's own error (2 pi 2^-c m)^2 / 6 and, at many slots, double-precision noise of EvalMod times q0 / (2 pi T) sqrt(s)
    assert np.abs(v - np.concatenate([z, z])).max() < (2.0 ** -17 if s <= 64 else 2.0 ** -15)
    rots = info["rotations"]
    assert s in rots and n // 4 in rots and all(0 < x < n // 2 for x in rots)
