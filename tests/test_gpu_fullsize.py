"""The configurations the headline numbers are quoted on, at their full parameters (ring 2^17), under pytest -m gpu:

  * EvalRotate / hoisted rotation / EvalMult at the DirectSort N = 1024 parameter set (depth 39, 41 + 10 limbs), bit-exact
    against the CPU oracle on identical limbs and keys (BASELINE.json config 2);
  * DirectSort N = 256 (depth 34, CompositeSign(3,4,2)) and N = 1024 (depth 39, (3,5,2)) with the reference's assertions
    (tests/DirectSortTest.cpp:128,169,172-179) - configs 3 and 5;
  * the HKC+21 k-way network k = 2, M = 10 (N = 1024) with bootstrapping (tests/k-way/KWaySort2Test.cpp:137-191) - config 5;
  * MEHP24 N = 256 lives in tests/test_gpu_mehp24.py - config 4.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEED = 1170


@pytest.fixture(scope="module")
def big():
    import fhe_sorting_b200 as F
    from oracle.oracle import Oracle
    ctx = F.Context(17, 39, 40, 60)
    orc = Oracle(17, 39, 40, 60, 3, threads=16)
    assert ctx.moduli == orc.moduli and (orc.L, orc.K) == (41, 10)
    keys = ctx.keygen(SEED)
    ctx.eval_mult_keygen(keys)
    ctx.eval_rotate_keygen(keys, [1, -3])
    s = orc.keygen_secret(SEED)
    return ctx, orc, keys, s


def _fresh(orc, s, rng, ell, slots=1024):
    vals = rng.uniform(-1, 1, slots)
    scale = orc.sf_big(0) if ell == orc.L else orc.sf(orc.L - ell)
    pk = orc.keygen_public(SEED, s)
    pt = orc.encode(vals, slots, scale, ell)
    c0, c1 = orc.encrypt(SEED, 5000 + ell, (pk[0][:ell], pk[1][:ell]), pt)
    return vals, scale, c0, c1


def test_ntt_every_modulus_class_ring17(big):
    ctx, orc, _, _ = big
    rng = np.random.default_rng(1)
    for qp in (0, 1, 17, orc.L - 1, orc.L, orc.L + orc.K - 1):     # 60-bit first, 40-bit scaling, extra, 60-bit special
        a = rng.integers(0, orc.moduli[qp], size=orc.n, dtype=np.uint64)
        f = ctx.raw_ntt(qp, a)
        assert np.array_equal(f, orc.ntt_fwd(qp, a)), "forward NTT, modulus %d" % qp
        assert np.array_equal(ctx.raw_ntt(qp, f, inverse=True), a), "inverse NTT, modulus %d" % qp


def test_modup_and_rescale_41_limbs(big):
    ctx, orc, _, _ = big
    rng = np.random.default_rng(2)
    for ell in (41, 28, 15):
        c = orc.random_poly(rng, ell)
        assert np.array_equal(ctx.raw_modup(c), orc.modup(c)), "ModUp at %d limbs" % ell
    a = orc.random_poly(rng, 41)
    assert np.array_equal(ctx.raw_rescale(a), orc.rescale(a))


def test_rotate_hoisted_mult_bit_exact_41_limbs(big):
    """The bench's workload: EvalRotate at 41 limbs of ring 2^17, every limb of both polynomials equal to the oracle's."""
    ctx, orc, keys, s = big
    rng = np.random.default_rng(3)
    for ell in (41, 20):
        vals, scale, c0, c1 = _fresh(orc, s, rng, ell)
        ct = ctx.ct_import(c0, c1, 2 if ell == orc.L else 1, scale, 1024)
        pre = ctx.fast_rotation_precompute(ct)
        for rot in (1, -3):
            g = orc.galois(rot)
            assert g == ctx.galois(rot)
            e0, e1 = orc.rotate(g, c0, c1, orc.keygen_rot(SEED, g, s))
            r0, r1 = ctx.ct_export(ctx.rotate(ct, rot))
            assert np.array_equal(r0, e0) and np.array_equal(r1, e1), "EvalRotate ell=%d rot=%d" % (ell, rot)
            h0, h1 = ctx.ct_export(ctx.fast_rotation(ct, rot, pre))
            assert np.array_equal(h0, e0) and np.array_equal(h1, e1), "EvalFastRotation ell=%d rot=%d" % (ell, rot)
        got = ctx.decrypt(keys, ctx.rotate(ct, 1), 1024)
        assert np.abs(got - np.roll(vals, -1)).max() < 1e-6
    evm = orc.keygen_mult(SEED, s)
    va, scale, a0, a1 = _fresh(orc, s, rng, 40)
    vb, _, b0, b1 = _fresh(orc, s, rng, 40)
    A, B = ctx.ct_import(a0, a1, 1, scale, 1024), ctx.ct_import(b0, b1, 1, scale, 1024)
    e0, e1 = orc.mult_relin(a0, a1, b0, b1, evm)
    r0, r1 = ctx.ct_export(ctx.mult(A, B))
    assert np.array_equal(r0, e0) and np.array_equal(r1, e1), "EvalMult at 40 limbs"


def test_mult_rescaled_one_division_41_limbs(big):
    """fheb200_eval_mult_rescaled at the full parameter set: level / scale of rescale(mult), coefficients within one rounding."""
    from test_gpu_scheme import _centered_coeff_diff
    ctx, orc, keys, s = big
    rng = np.random.default_rng(8)
    va, vb = rng.uniform(-1, 1, 1024), rng.uniform(-1, 1, 1024)
    a = ctx.encrypt(keys, ctx.make_plaintext(va, 1, 0, 1024))
    b = ctx.encrypt(keys, ctx.make_plaintext(vb, 1, 0, 1024))
    two, one = ctx.rescale(ctx.mult(a, b)), ctx.mult_rescaled(a, b)
    assert one.info() == two.info() and one.info()["limbs"] == 39
    assert np.abs(ctx.decrypt(keys, one, 1024) - va * vb).max() < 1e-5
    assert np.abs(ctx.decrypt(keys, one, 1024) - ctx.decrypt(keys, two, 1024)).max() < 1e-9
    x0, _ = ctx.ct_export(one)
    y0, _ = ctx.ct_export(two)
    worst, same = _centered_coeff_diff(ctx, x0[:3], y0[:3])
    assert worst <= 1 and same, (worst, same)


@pytest.mark.parametrize("n,depth", [(256, 34), (1024, 39)])
def test_direct_sort_full_parameters(n, depth):
    """tests/DirectSortTest.cpp:172-179 sizes 256 and 1024: ring 2^17, the per-N depth table of src/sort_algo.h:87-201."""
    from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
    s = DirectSortSession(n, 17, seed=SEED + n)
    assert s.mult_depth == depth
    x = uniform_gap_input(n)
    out, ms, level = s.sort(x)
    err = np.abs(out - np.sort(x)).max()
    print("DirectSort N=%d ring 2^17: %.0f ms, max error %.2e, level %d, %d rotation keys" % (n, ms, err, level, s.num_rotation_keys))
    assert err < 0.01                                      # tests/DirectSortTest.cpp:169
    assert level == s.mult_depth                           # tests/DirectSortTest.cpp:128
    ranks, _, _ = s.rank(x)
    assert np.abs(ranks - np.argsort(np.argsort(x))).max() < 0.01
    s.close()


def test_kway_k2_m10_full_parameters():
    """tests/k-way/KWaySort2Test.cpp:137-191: N = 1024, ring 2^17, depth 40, 59-bit scaling, bootstrapping {5,5}."""
    from fhe_sorting_b200 import sort as S
    s = S.KWaySortSession(2, 10, log_ring_dim=17, seed=SEED + 2)
    assert s.mult_depth == 40
    x = np.random.default_rng(1024).permutation(1024) * ((1 - 1e-8) / 1024)
    out, ms, level = s.sort(x)
    err = np.abs(out - np.sort(x)).max()
    print("k-way k=2 N=1024: %.0f ms, %d guard bootstraps, max error %.2e" % (ms, s.bootstraps(), err))
    assert err < 0.01                                      # KWaySort2Test.cpp:190
    s.close()
