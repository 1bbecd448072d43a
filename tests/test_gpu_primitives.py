"""Bit-exact parity of every integer RNS kernel against the CPU oracle (through the C ABI), -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEED = 0xC0FFEE


@pytest.fixture(scope="module", params=[(10, 6, 40), (12, 9, 50), (13, 5, 59), (14, 4, 40)], ids=lambda p: "n2^%d_d%d_s%d" % p)
def pair(request):
    import fhe_sorting_b200 as F
    from oracle.oracle import Oracle
    logn, depth, sb = request.param
    ctx = F.Context(logn, depth, sb, 60)
    orc = Oracle(logn, depth, sb, 60, 3, threads=4)
    assert ctx.moduli == orc.moduli
    keys = ctx.keygen(SEED)
    ctx.eval_mult_keygen(keys)
    ctx.eval_rotate_keygen(keys, [1, -1, 3])
    s = orc.keygen_secret(SEED)
    return ctx, orc, keys, s


def test_ntt_forward_inverse(pair):
    ctx, orc, _, _ = pair
    rng = np.random.default_rng(1)
    for qp in sorted({0, 1, orc.L - 1, orc.L, orc.L + orc.K - 1}):
        a = rng.integers(0, orc.moduli[qp], size=orc.n, dtype=np.uint64)
        f = ctx.raw_ntt(qp, a)
        assert np.array_equal(f, orc.ntt_fwd(qp, a)), "forward NTT mismatch at modulus %d" % qp
        assert np.array_equal(ctx.raw_ntt(qp, f, inverse=True), a)
        assert np.array_equal(ctx.raw_ntt(qp, a, inverse=True), orc.ntt_inv(qp, a))
    # edge values: zeros, q-1 everywhere
    q = orc.moduli[0]
    for a in (np.zeros(orc.n, dtype=np.uint64), np.full(orc.n, q - 1, dtype=np.uint64)):
        assert np.array_equal(ctx.raw_ntt(0, a), orc.ntt_fwd(0, a))


def test_automorphism(pair):
    ctx, orc, _, _ = pair
    rng = np.random.default_rng(2)
    a = orc.random_poly(rng, 3)
    for rot in (1, -1, 3, 7, orc.n // 4):
        g = orc.galois(rot)
        assert g == ctx.galois(rot)
        assert np.array_equal(ctx.raw_automorphism(g, a), orc.automorphism_eval(g, a))


def test_modup_all_levels(pair):
    ctx, orc, _, _ = pair
    rng = np.random.default_rng(3)
    for ell in sorted({orc.L, orc.L - 1, orc.alpha + 1, orc.alpha, 2, 1}):
        c = orc.random_poly(rng, ell)
        assert np.array_equal(ctx.raw_modup(c), orc.modup(c)), "ModUp mismatch at %d limbs" % ell


def test_rescale(pair):
    ctx, orc, _, _ = pair
    rng = np.random.default_rng(4)
    for ell in (orc.L, 3, 2):
        a = orc.random_poly(rng, ell)
        assert np.array_equal(ctx.raw_rescale(a), orc.rescale(a))


def test_keys_match_oracle(pair):
    ctx, orc, keys, s = pair
    assert np.array_equal(ctx.keys_export(keys, 0), s)
    b, a = orc.keygen_public(SEED, s)
    assert np.array_equal(ctx.keys_export(keys, 2), a)
    assert np.array_equal(ctx.keys_export(keys, 1), b)
    assert np.array_equal(ctx.evk_export(0), orc.keygen_mult(SEED, s))
    g = orc.galois(3)
    assert np.array_equal(ctx.evk_export(g), orc.keygen_rot(SEED, g, s))


def _fresh(ctx, orc, s, rng, ell, slots=8, level_scale=None):
    vals = rng.uniform(-1, 1, slots)
    scale = orc.sf_big(0) if ell == orc.L else orc.sf(orc.L - ell)
    pk = orc.keygen_public(SEED, s)
    pt = orc.encode(vals, slots, scale, ell)
    c0, c1 = orc.encrypt(SEED, 1000 + ell, (pk[0][:ell], pk[1][:ell]), pt)
    return vals, scale, c0, c1


def test_rotate_and_hoisted_bit_exact(pair):
    ctx, orc, keys, s = pair
    rng = np.random.default_rng(5)
    for ell in sorted({orc.L, orc.alpha, 2}):
        vals, scale, c0, c1 = _fresh(ctx, orc, s, rng, ell)
        ct = ctx.ct_import(c0, c1, 2 if ell == orc.L else 1, scale, 8)
        for rot in (1, -1, 3):
            g = orc.galois(rot)
            evk = orc.keygen_rot(SEED, g, s)
            e0, e1 = orc.rotate(g, c0, c1, evk)
            r0, r1 = ctx.ct_export(ctx.rotate(ct, rot))
            assert np.array_equal(r0, e0) and np.array_equal(r1, e1), "EvalRotate mismatch ell=%d rot=%d" % (ell, rot)
            pre = ctx.fast_rotation_precompute(ct)
            h0, h1 = ctx.ct_export(ctx.fast_rotation(ct, rot, pre))
            assert np.array_equal(h0, e0) and np.array_equal(h1, e1)
        got = ctx.decrypt(keys, ctx.rotate(ct, 1), 8)
        assert np.abs(got - np.roll(vals, -1)).max() < 1e-6


def test_mult_relin_bit_exact(pair):
    ctx, orc, keys, s = pair
    rng = np.random.default_rng(6)
    evm = orc.keygen_mult(SEED, s)
    for ell in sorted({orc.L - 1, 2}):
        va, scale, a0, a1 = _fresh(ctx, orc, s, rng, ell)
        vb, _, b0, b1 = _fresh(ctx, orc, s, rng, ell)
        A = ctx.ct_import(a0, a1, 1, scale, 8)
        B = ctx.ct_import(b0, b1, 1, scale, 8)
        e0, e1 = orc.mult_relin(a0, a1, b0, b1, evm)
        r0, r1 = ctx.ct_export(ctx.mult(A, B))
        assert np.array_equal(r0, e0) and np.array_equal(r1, e1), "EvalMult mismatch at %d limbs" % ell
        s0, s1 = orc.mult_relin(a0, a1, a0, a1, evm)
        q0, q1 = ctx.ct_export(ctx.square(A))
        assert np.array_equal(q0, s0) and np.array_equal(q1, s1)


def test_encode_encrypt_match_oracle(pair):
    ctx, orc, keys, s = pair
    rng = np.random.default_rng(7)
    vals = rng.uniform(-1, 1, 16)
    pt = ctx.make_plaintext(vals, 1, 0, 16)
    mine = ctx.pt_export(pt, orc.L)
    # the encoder is floating point (special FFT, scale ~2^100): compare through the oracle's decoder, not bit for bit
    dec = orc.decrypt(mine, np.zeros_like(mine), s, 16, orc.sf_big(0))
    assert np.abs(dec - vals).max() < 1e-9, "CKKS encode does not decode under the oracle"
    got = ctx.decrypt(keys, ctx.encrypt(keys, pt), 16)
    assert np.abs(got - vals).max() < 1e-9


def test_async_import_export_pipeline(pair):
    """fheb200_ct_import_async / _ready / _export_async: a pipelined sequence of work items over pinned host buffers
    gives the limbs of the synchronous path (and so of the oracle), whatever overlaps."""
    import torch
    ctx, orc, keys, s = pair
    rng = np.random.default_rng(11)
    ell = orc.L
    items = [_fresh(ctx, orc, s, rng, ell) for _ in range(4)]
    pins = [torch.empty((ell, orc.n), dtype=torch.int64).pin_memory() for _ in range(4 * len(items))]
    bufs = [p.numpy().view(np.uint64) for p in pins]
    h_in = [(bufs[4 * k], bufs[4 * k + 1]) for k in range(len(items))]
    h_out = [(bufs[4 * k + 2], bufs[4 * k + 3]) for k in range(len(items))]
    for (vals, scale, c0, c1), (i0, i1) in zip(items, h_in):
        i0[:], i1[:] = c0, c1
    want = []
    for vals, scale, c0, c1 in items:
        ct = ctx.ct_import(c0, c1, 2, scale, 8)
        want.append(ctx.ct_export(ctx.rotate(ctx.rotate(ct, 1), 3)))
    g1, g3 = orc.galois(1), orc.galois(3)
    e = orc.rotate(g1, items[0][2], items[0][3], orc.keygen_rot(SEED, g1, s))
    e = orc.rotate(g3, e[0], e[1], orc.keygen_rot(SEED, g3, s))
    assert np.array_equal(want[0][0], e[0]) and np.array_equal(want[0][1], e[1])
    nxt = ctx.ct_import_async(h_in[0][0], h_in[0][1], 2, items[0][1], 8)
    for k in range(len(items)):
        cur = nxt
        if k + 1 < len(items):
            nxt = ctx.ct_import_async(h_in[k + 1][0], h_in[k + 1][1], 2, items[k + 1][1], 8)
        ctx.ct_ready(cur)
        r = ctx.rotate(ctx.rotate(cur, 1), 3)
        ctx.ct_export_async(r, h_out[k])
        del cur, r  # the engine keeps the buffers alive until the copies have run
    ctx.sync()
    for k in range(len(items)):
        assert np.array_equal(h_out[k][0], want[k][0]) and np.array_equal(h_out[k][1], want[k][1]), "item %d" % k
    # an import that is never waited for explicitly is still complete after sync()
    lone = ctx.ct_import_async(h_in[1][0], h_in[1][1], 2, items[1][1], 8)
    ctx.sync()
    back = ctx.ct_export(lone)
    assert np.array_equal(back[0], items[1][2]) and np.array_equal(back[1], items[1][3])


def test_lincomb_kernel_variants_are_bit_identical():
    """The Paterson-Stockmeyer leaf-sum kernel has two implementations (FHEB200_LINCOMB=1: bounds-checked register loads,
    5: branch-free padded groups through a cp.async ring).  Same partial sums in the same order: identical output
    checksums on 63 inputs -> 4 outputs at three limb counts (fheb200_bench_kernel("lincomb") prints them)."""
    import os
    import re
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sums = {}
    for v in ("1", "5"):
        o = subprocess.run([sys.executable, os.path.join(root, "tools", "lincomb_bench.py")], capture_output=True, text=True,
                           env=dict(os.environ, FHEB200_LINCOMB=v), timeout=300)
        assert o.returncode == 0, o.stderr[-2000:]
        sums[v] = re.findall(r"lincomb checksum ([0-9a-f]{16})", o.stderr)
        assert len(sums[v]) == 3
    assert sums["1"] == sums["5"]
