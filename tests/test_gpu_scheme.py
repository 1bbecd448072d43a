"""Scheme-level behaviour through the C ABI: FLEXIBLEAUTOEXT level/scale bookkeeping, plaintext ops, Chebyshev /
polynomial evaluation, and the reference's own known-answer vectors (tests/SignTest.cpp, CompareTest.cpp,
RotationTest.cpp of oksuman/FHE-Sorting).  Decrypted-value tolerances are the reference's."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import fhe_sorting_b200 as F
    ctx = F.Context(12, 30, 50, 60, batch=8)          # tests/SignTest.cpp:24-36 (ring 2^12, depth 30, scale 50)
    keys = ctx.keygen(11)
    ctx.eval_mult_keygen(keys)
    ctx.eval_rotate_keygen(keys, [1, 2, -1, -2, 4])
    return ctx, keys


def enc(ctx, keys, v, slots=8):
    return ctx.encrypt(keys, ctx.make_plaintext(v, 1, 0, slots))


def test_fresh_level_and_roundtrip(env):
    ctx, keys = env
    v = np.array([0.5, -0.25, 0.125, 1, -1, 0, 0.75, -0.6])
    c = enc(ctx, keys, v)
    i = c.info()
    assert i["level"] == 0 and i["nsd"] == 2 and i["limbs"] == ctx.L and i["slots"] == 8
    assert np.abs(ctx.decrypt(keys, c, 8) - v).max() < 1e-9


def test_levels_follow_flexibleautoext(env):
    ctx, keys = env
    v = np.linspace(-0.9, 0.9, 8)
    c = enc(ctx, keys, v)
    m1 = ctx.mult(c, c)                 # first mult rescales the extra modulus away: level 1, degree 2
    assert (m1.level, m1.info()["nsd"]) == (1, 2)
    m2 = ctx.mult(m1, 0.5)              # scalar mult counts as a multiplication
    assert (m2.level, m2.info()["nsd"]) == (2, 2)
    m3 = ctx.add(m2, 1.0)
    assert m3.level == 2
    m4 = ctx.mult(m3, c)                # operands at different levels are aligned automatically
    assert m4.level == 3
    want = (v * v * 0.5 + 1.0) * v
    assert np.abs(ctx.decrypt(keys, m4, 8) - want).max() < 1e-7


def test_add_sub_mixed_levels(env):
    ctx, keys = env
    a = np.linspace(-1, 1, 8)
    b = np.linspace(0.1, 0.8, 8)
    ca, cb = enc(ctx, keys, a), enc(ctx, keys, b)
    deep = ctx.mult(ctx.mult(ca, ca), ca)
    for other in (cb, ctx.mult(cb, 1.0), ctx.mult(cb, cb)):
        bo = ctx.decrypt(keys, other, 8)
        assert np.abs(ctx.decrypt(keys, ctx.add(deep, other), 8) - (a ** 3 + bo)).max() < 1e-7
        assert np.abs(ctx.decrypt(keys, ctx.sub(other, deep), 8) - (bo - a ** 3)).max() < 1e-7
    assert np.abs(ctx.decrypt(keys, ctx.negate(deep), 8) + a ** 3).max() < 1e-7
    assert np.abs(ctx.decrypt(keys, ctx.add_many([ca, cb, deep]), 8) - (a + b + a ** 3)).max() < 1e-7


def test_plaintext_ops(env):
    ctx, keys = env
    a = np.linspace(-1, 1, 8)
    m = np.array([1, 0, 1, 0, 0.5, 0, 2, -1.0])
    ca = enc(ctx, keys, a)
    pm = ctx.make_plaintext(m, 1, 0, 8)
    r = ctx.mult(ca, pm)
    assert r.level == 1
    assert np.abs(ctx.decrypt(keys, r, 8) - a * m).max() < 1e-8
    deep = ctx.mult(ca, ca)
    assert np.abs(ctx.decrypt(keys, ctx.mult(deep, pm), 8) - a * a * m).max() < 1e-8
    assert np.abs(ctx.decrypt(keys, ctx.add(deep, pm), 8) - (a * a + m)).max() < 1e-8
    assert np.abs(ctx.decrypt(keys, ctx.sub(deep, pm), 8) - (a * a - m)).max() < 1e-8
    pl = ctx.make_plaintext(m, 1, deep.level, 8)      # MakeCKKSPackedPlaintext(vec, 1, level, nullptr, slots)
    assert np.abs(ctx.decrypt(keys, ctx.sub(deep, pl), 8) - (a * a - m)).max() < 1e-8


def test_rotation_known_answers(env):
    # tests/RotationTest.cpp:84-94 — positive index rotates left
    ctx, keys = env
    v = np.arange(1, 9, dtype=float)
    c = enc(ctx, keys, v)
    for r in (1, 2, -1, -2, 4):
        assert np.abs(ctx.decrypt(keys, ctx.rotate(c, r), 8) - np.roll(v, -r)).max() < 1e-6
    back = ctx.rotate(ctx.rotate(c, 2), -2)
    assert np.abs(ctx.decrypt(keys, back, 8) - v).max() < 1e-5   # RotationTest.cpp:126
    with pytest.raises(Exception):
        ctx.rotate(c, 3)                                         # no key for index 3


def test_set_slots_reinterprets_packing(env):
    # src/sort_algo.h:429,501: an N-slot ciphertext is a valid ciphertext of any multiple of N slots
    ctx, keys = env
    v = np.arange(1, 9, dtype=float)
    c = enc(ctx, keys, v)
    c.set_slots(32)
    got = ctx.decrypt(keys, c, 32)
    assert np.abs(got - np.tile(v, 4)).max() < 1e-8


def g3(x):
    return (4589 * x - 16577 * x ** 3 + 25614 * x ** 5 - 12860 * x ** 7) / 1024


def f3(x):
    return (35 * x - 35 * x ** 3 + 21 * x ** 5 - 5 * x ** 7) / 16


def composite_sign_n3(ctx, x, dg, df):
    """src/sign.cpp:15-59,160-185 written against the binding (same operation order)."""
    def poly(x, c1, c3, c5, c7):
        x2 = ctx.square(x)
        x4 = ctx.square(x2)
        y = ctx.mult(x, c1)
        y = ctx.add(y, ctx.mult(ctx.mult(x, c3), x2))
        c5x, c7x = ctx.mult(x, c5), ctx.mult(x, c7)
        c7x3 = ctx.mult(c7x, x2)
        return ctx.add(y, ctx.mult(ctx.add(c5x, c7x3), x4))
    y = x
    for _ in range(dg):
        y = poly(y, 4589 / 1024, -16577 / 1024, 25614 / 1024, -12860 / 1024)
    for _ in range(df):
        y = poly(y, 35 / 16, -35 / 16, 21 / 16, -5 / 16)
    return y


def test_composite_sign_known_answer(env):
    # tests/SignTest.cpp:41-70: compositeSign<3>, |err| < 0.1 on +-1 ; each g/f consumes exactly 3 levels
    ctx, keys = env
    v = np.array([-0.9, -0.5, -0.2, -0.05, 0.05, 0.2, 0.5, 0.9])
    c = enc(ctx, keys, v)
    y = composite_sign_n3(ctx, c, 3, 3)
    assert y.level == 18
    got = ctx.decrypt(keys, y, 8)
    assert np.abs(got - np.sign(v)).max() < 0.1
    ref = v.copy()
    for _ in range(3):
        ref = g3(ref)
    for _ in range(3):
        ref = f3(ref)
    assert np.abs(got - ref).max() < 1e-4


def test_compare_known_answer(env):
    # tests/CompareTest.cpp:43-63 {1,5,3,4} vs {2,4,3,3} (scaled into [-1,1]) -> {0,1,0.5,1} +- 0.1
    ctx, keys = env
    a = np.array([1, 5, 3, 4, 0, 0, 0, 0]) / 8
    b = np.array([2, 4, 3, 3, 0, 0, 0, 0]) / 8
    d = ctx.sub(enc(ctx, keys, a), enc(ctx, keys, b))
    comp = ctx.mult(ctx.add(composite_sign_n3(ctx, d, 3, 3), 1.0), 0.5)   # src/comparison.cpp:13-19
    got = ctx.decrypt(keys, comp, 4)
    assert np.abs(got - np.array([0, 1, 0.5, 1])).max() < 0.1


def cheb_eval(coeffs, x):
    from numpy.polynomial import chebyshev as Ch
    c = np.array(coeffs, dtype=float).copy()
    c[0] /= 2
    return Ch.chebval(x, c)


@pytest.mark.parametrize("deg", [3, 7, 27, 59, 126, 200])
def test_chebyshev_series(env, deg):
    ctx, keys = env
    rng = np.random.default_rng(deg)
    coeffs = rng.uniform(-1, 1, deg + 1) / np.arange(1, deg + 2)
    v = np.linspace(-0.95, 0.95, 8)
    c = enc(ctx, keys, v)
    y = ctx.chebyshev(c, coeffs, -1.0, 1.0)
    assert np.abs(ctx.decrypt(keys, y, 8) - cheb_eval(coeffs, v)).max() < 1e-6
    # levels consumed = min over (k, m) with k(2^m-1) > deg of ceil(log2 k) + m  (OpenFHE's depth table)
    if deg >= 5:
        want = min(int(np.ceil(np.log2(k))) + m for k in range(2, deg + 2) for m in range(1, 12) if k * (2 ** m - 1) > deg)
        assert y.level == want, (deg, y.level, want)


def test_chebyshev_on_interval_and_poly(env):
    ctx, keys = env
    v = np.linspace(0.5, 3.5, 8)
    c = enc(ctx, keys, v)
    from numpy.polynomial import chebyshev as Ch
    coeffs = np.array([2 * x for x in Ch.chebinterpolate(lambda t: np.sin(2 * t + 2), 20)])
    coeffs[0] = coeffs[0]        # series convention: c0/2
    y = ctx.chebyshev(c, coeffs, 0.0, 4.0)
    u = (2 * v - 4.0) / 4.0
    assert np.abs(ctx.decrypt(keys, y, 8) - cheb_eval(coeffs, u)).max() < 1e-6
    p = [0.0, 35 / 16, 0.0, -35 / 16, 0.0, 21 / 16, 0.0, -5 / 16]     # src/mehp24/mehp24_utils.cpp:246-261
    w = np.linspace(-1, 1, 8)
    z = ctx.poly(enc(ctx, keys, w), p)
    assert np.abs(ctx.decrypt(keys, z, 8) - f3(w)).max() < 1e-7
    assert z.level == 4


def test_rotated_plaintext_is_a_galois_permutation_of_the_encoding(env):
    """fheb200_pt_rotate: slot i of the result is slot i + r of the plaintext, for sparse (8 of 2048) and full packing,
    positive and negative r; used as a mask it equals the directly encoded rotated vector."""
    import fhe_sorting_b200 as F
    ctx, keys = env
    v = np.array([0.5, -0.25, 0.125, 1.0, -1.0, 0.0, 0.75, -0.6])
    x = np.array([1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0, 8.0]) / 8
    ct = enc(ctx, keys, x)
    base = ctx.make_plaintext(v, 1, 0, 8)
    for r in (1, 3, -2, 8, 13):
        rp = ctx.rotate_plaintext(base, r)
        want = np.roll(v, -r)
        assert np.abs(ctx.decrypt(keys, ctx.encrypt(keys, rp), 8) - want).max() < 1e-9
        direct = ctx.mult(ct, ctx.make_plaintext(want, 1, 0, 8))
        via = ctx.mult(ct, rp)
        assert np.abs(ctx.decrypt(keys, via, 8) - x * want).max() < 1e-9
        assert np.abs(ctx.decrypt(keys, via, 8) - ctx.decrypt(keys, direct, 8)).max() < 1e-9
    full = F.Context(12, 4, 40, 60, batch=2048)
    k2 = full.keygen(3)
    w = np.random.default_rng(2).uniform(-1, 1, 2048)
    b2 = full.make_plaintext(w, 1, 0, 2048)
    for r in (5, -700):
        got = full.decrypt(k2, full.encrypt(k2, full.rotate_plaintext(b2, r)), 2048)
        assert np.abs(got - np.roll(w, -r)).max() < 1e-7


def _centered_coeff_diff(ctx, x, y):
    """max over limbs and coefficients of |x - y| (centred, coefficient domain) and whether the difference is the same small
    integer in every limb (x, y: [ell][n] evaluation-domain limbs)"""
    worst, same = 0, True
    first = None
    for i in range(x.shape[0]):
        q = ctx.moduli[i]
        d = (ctx.raw_ntt(i, x[i], inverse=True).astype(object) - ctx.raw_ntt(i, y[i], inverse=True).astype(object)) % q
        d = np.array([int(v) - q if int(v) > q // 2 else int(v) for v in d], dtype=np.int64)
        worst = max(worst, int(np.abs(d).max()))
        if first is None:
            first = d
        same = same and np.array_equal(d, first)
    return worst, same


def test_mult_rescaled_is_one_division(env):
    """relinearise + rescale as one ModDown by P * q_last: same level, degree and scale as rescale(mult(a, b)); in the
    coefficient domain the two results differ by the same small integer in every limb (one rounding instead of two)."""
    ctx, keys = env
    rng = np.random.default_rng(5)
    va, vb = rng.uniform(-1, 1, 8), rng.uniform(-1, 1, 8)
    a, b = enc(ctx, keys, va), enc(ctx, keys, vb)
    for _ in range(3):                                   # at several levels
        two = ctx.rescale(ctx.mult(a, b))
        one = ctx.mult_rescaled(a, b)
        assert one.info() == two.info()
        assert np.abs(ctx.decrypt(keys, one, 8) - va * vb).max() < 1e-9
        x0, x1 = ctx.ct_export(one)
        y0, y1 = ctx.ct_export(two)
        for x, y in ((x0, y0), (x1, y1)):
            worst, same = _centered_coeff_diff(ctx, x, y)
            assert worst <= 1 and same, (worst, same)   # exact (centred) lift: at most the double rounding of the two-step form
        a, vb = one, vb
        va = va * vb
