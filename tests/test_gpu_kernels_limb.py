"""Limb-level parity of the kernels that round 1 only compared with themselves: each one against the CPU oracle or
against exact Python integers on the same limbs (-m gpu, through the C ABI's raw entry points).

  lc_tc_ring / lc_tc / ew_lincomb   Paterson-Stockmeyer leaf sums    vs  sum_t w x mod q in Python integers
  ew_mul_pt (+ accumulate)          ct x plaintext                    vs  orc_mul / orc_add
  ew_tensor / ew_tensor_square      EvalMult before relinearisation   vs  orc_tensor
  ks_inner, ModDown pipeline        the two halves of a key switch    vs  orc_inner_product / orc_moddown
  ks_dh_accumulate                  double-hoisted accumulation in QP vs  Python integers
  fheb200_eval_mult_plain_add       acc + ct * pt in one pass         vs  EvalMult(ct, pt) then EvalAdd, limb for limb
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEED = 0xBEEF


def _ctx(logn, depth, sb):
    import fhe_sorting_b200 as F
    from oracle.oracle import Oracle
    ctx = F.Context(logn, depth, sb, 60)
    orc = Oracle(logn, depth, sb, 60, 3, threads=8)
    assert ctx.moduli == orc.moduli
    return ctx, orc


@pytest.fixture(scope="module", params=[(12, 39, 40), (12, 12, 59)], ids=["40bit_scaling", "59bit_scaling"])
def pair(request):
    return _ctx(*request.param)


def _exact_lincomb(inputs, weights, moduli, ell, rows):
    """out[g][limb][rows] with Python integers (object arrays): sum_t w[g][t][limb % ell] * in[t][limb] mod q"""
    count, limbs2 = inputs.shape[0], inputs.shape[1]
    G = weights.shape[0]
    out = np.zeros((G, limbs2, len(rows)), dtype=np.uint64)
    x = inputs[:, :, rows].astype(object)
    for g in range(G):
        for li in range(limbs2):
            q = moduli[li % ell]
            acc = np.zeros(len(rows), dtype=object)
            for t in range(count):
                acc = acc + x[t, li] * int(weights[g, t, li % ell])
            out[g, li] = np.array([int(v) % q for v in acc], dtype=np.uint64)
    return out


@pytest.mark.parametrize("ell", [4, 14, 41])
@pytest.mark.parametrize("count,G", [(63, 8), (9, 3), (16, 8)])
def test_leaf_sums_all_variants(pair, ell, count, G):
    """k = 63 inputs, 8 outputs is the degree-6510 leaf pass of DirectSort N = 1024 (src/sort_algo.h:557)."""
    ctx, orc = pair
    if ell > orc.L:
        pytest.skip("chain shorter than %d limbs" % ell)
    rng = np.random.default_rng(ell * 1000 + count)
    npoly = 2
    inputs = np.stack([np.concatenate([orc.random_poly(rng, ell) for _ in range(npoly)]) for _ in range(count)])
    weights = np.zeros((G, count, ell), dtype=np.uint64)
    for i in range(ell):
        weights[:, :, i] = rng.integers(0, orc.moduli[i], size=(G, count), dtype=np.uint64)
    weights[0, 0, :] = 0                                   # zero and maximal weights
    weights[G - 1, count - 1, :] = np.array(orc.moduli[:ell], dtype=np.uint64) - np.uint64(1)
    inputs[0, :, :7] = 0
    for i in range(ell):                                   # maximal residues
        inputs[1, i, :9] = np.uint64(orc.moduli[i] - 1)
    rows = np.concatenate([np.arange(40), rng.integers(0, orc.n, 88), [orc.n - 1]])
    want = _exact_lincomb(inputs, weights, orc.moduli, ell, rows)
    outs = {}
    for variant, name in ((0, "lc_tc_ring"), (1, "lc_tc"), (2, "ew_lincomb"), (3, "lc_tc_t5")):
        g_lim = min(G, 4) if variant == 2 else G
        outs[name] = ctx.raw_lincomb(inputs, weights[:g_lim], variant)
        assert np.array_equal(outs[name][:, :, rows], want[:g_lim]), name
    # and everywhere else the three kernels agree with each other
    assert np.array_equal(outs["lc_tc_ring"], outs["lc_tc"])
    assert np.array_equal(outs["lc_tc_ring"], outs["lc_tc_t5"]), "tcgen05 leaf sums differ from the mma.sync ones"
    assert np.array_equal(outs["lc_tc_ring"][:outs["ew_lincomb"].shape[0]], outs["ew_lincomb"])


@pytest.mark.parametrize("ell", [2, 14, 41])
def test_mul_plaintext_and_tensor(pair, ell):
    ctx, orc = pair
    if ell > orc.L:
        pytest.skip("chain shorter than %d limbs" % ell)
    rng = np.random.default_rng(ell)
    a0, a1, b0, b1, pt, s0, s1 = (orc.random_poly(rng, ell) for _ in range(7))
    o0, o1 = ctx.raw_mul_pt(a0, a1, pt)
    assert np.array_equal(o0, orc.mul(a0, pt)) and np.array_equal(o1, orc.mul(a1, pt))
    o0, o1 = ctx.raw_mul_pt(a0, a1, pt, acc=(s0, s1))
    assert np.array_equal(o0, orc.add(s0, orc.mul(a0, pt))) and np.array_equal(o1, orc.add(s1, orc.mul(a1, pt)))
    d = ctx.raw_tensor(a0, a1, b0, b1)
    w = orc.tensor(a0, a1, b0, b1)
    assert all(np.array_equal(x, y) for x, y in zip(d, w)), "ew_tensor"
    d = ctx.raw_tensor(a0, a1)
    w = orc.tensor(a0, a1, a0, a1)
    assert all(np.array_equal(x, y) for x, y in zip(d, w)), "ew_tensor_square"


def test_inner_product_and_moddown(pair):
    ctx, orc = pair
    keys = ctx.keygen(SEED)
    ctx.eval_mult_keygen(keys)
    ctx.eval_rotate_keygen(keys, [5])
    s = orc.keygen_secret(SEED)
    g = orc.galois(5)
    rng = np.random.default_rng(17)
    for ell in sorted({orc.L, orc.alpha + 1, 2}):
        W = ell + orc.K
        digits = np.zeros((orc.beta(ell), W, orc.n), dtype=np.uint64)
        for t in range(W):
            q = orc.moduli[t if t < ell else orc.L + (t - ell)]
            digits[:, t, :] = rng.integers(0, q, size=(digits.shape[0], orc.n), dtype=np.uint64)
        for gal, evk in ((0, orc.keygen_mult(SEED, s)), (g, orc.keygen_rot(SEED, g, s))):
            acc = ctx.raw_inner_product(ell, gal, digits)
            w0, w1 = orc.inner_product(ell, digits, evk)
            assert np.array_equal(acc[0], w0) and np.array_equal(acc[1], w1), "ks_inner at %d limbs" % ell
        o0, o1 = ctx.raw_moddown(ell, acc)
        assert np.array_equal(o0, orc.moddown(ell, acc[0])) and np.array_equal(o1, orc.moddown(ell, acc[1])), \
            "ModDown at %d limbs" % ell


def test_double_hoisting_accumulate(pair):
    """S_b[t][p] += pt[t][p] * v_b[t][sigma_g(p)] over the extended basis; v_0 = acc_0 + cP_0, v_1 = acc_1 (rotated term) or
    v_b = cP_b on the Q limbs only (acc null: the unrotated term)."""
    ctx, orc = pair
    rng = np.random.default_rng(23)
    n = orc.n
    for ell in sorted({orc.L, 3}):
        W = ell + orc.K
        mods = [orc.moduli[t if t < ell else orc.L + (t - ell)] for t in range(W)]

        def rnd(limbs, lead=()):
            out = np.zeros(tuple(lead) + (limbs, n), dtype=np.uint64)
            for t in range(limbs):
                out[..., t, :] = rng.integers(0, mods[t], size=tuple(lead) + (n,), dtype=np.uint64)
            return out
        acc, cP, pt, S = rnd(W, (2,)), rnd(ell, (2,)), rnd(W), rnd(W, (2,))
        for g in (1, orc.galois(3), orc.galois(-7)):
            perm = np.arange(n, dtype=np.uint64).reshape(1, n)
            src = orc.automorphism_eval(g, perm)[0].astype(np.int64) if g != 1 else np.arange(n)
            for use_acc in (True, False):
                got = ctx.raw_dh_accumulate(ell, acc if use_acc else None, cP, pt, S, g)
                want = S.copy()
                for t in range(W):
                    q = mods[t]
                    if not use_acc and t >= ell:
                        continue
                    for b in range(2):
                        v = acc[b, t][src].astype(object) if use_acc else np.zeros(n, dtype=object)
                        if t < ell and (b == 0 or not use_acc):
                            v = v + cP[b, t][src].astype(object)
                        r = (S[b, t].astype(object) + v * pt[t].astype(object)) % q
                        want[b, t] = r.astype(np.uint64)
                assert np.array_equal(got, want), "ks_dh_accumulate ell=%d g=%d acc=%s" % (ell, g, use_acc)


def test_mult_plain_add_equals_mult_then_add(pair):
    """fheb200_eval_mult_plain_add (the sort drivers' mask-and-accumulate loops) limb for limb."""
    import ctypes as C
    import fhe_sorting_b200 as F
    from fhe_sorting_b200.lib import _ck
    _lib = F.load()
    ctx, orc = pair
    keys = ctx.keygen(SEED + 1)
    rng = np.random.default_rng(5)
    slots = 64
    va, vb, vm = rng.uniform(-1, 1, slots), rng.uniform(-1, 1, slots), (rng.uniform(0, 1, slots) > 0.5).astype(float)
    a = ctx.encrypt(keys, ctx.make_plaintext(va, 1, 0, slots))
    pm = ctx.make_plaintext(vm, 1, 0, slots)
    for depth_a in (0, 2):
        x = a
        for _ in range(depth_a):
            x = ctx.mult(x, 1.0)
        acc = ctx.mult(x, ctx.make_plaintext(vb, 1, 0, slots))   # sits where x * pm lands
        two = ctx.add(acc, ctx.mult(x, pm))
        h = C.c_void_p()
        _ck(_lib.fheb200_eval_mult_plain_add(ctx.h, acc.h, x.h, pm.h, C.byref(h)))
        one = F.Ciphertext(h)
        assert one.info() == two.info()
        o, t = ctx.ct_export(one), ctx.ct_export(two)
        assert np.array_equal(o[0], t[0]) and np.array_equal(o[1], t[1])
        got = ctx.decrypt(keys, one, slots)
        assert np.abs(got - (va * vb + va * vm)).max() < 1e-7
