"""CKKS bootstrapping on the B200 engine, checked step by step against what each step must hold in the clear
(EvalBootstrapSetup / EvalBootstrapKeyGen / EvalBootstrap, reference call sites tests/k-way/KWaySort2Test.cpp:42-44,
src/k-way/EvalUtils.cpp:57-86).  Parameters as the k-way fixture: first modulus 60 bits, scaling 59 bits."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from test_bootstrap_plan import bitrev, special_matrix  # noqa: E402

CORRECTION = 10  # BootPlan::correction default (FHEB200_BOOT_CORRECTION)


def make(logn, depth, slots, budget, seed=5):
    import fhe_sorting_b200 as F
    ctx = F.Context(logn, depth, 59, 60, batch=slots)
    keys = ctx.keygen(seed)
    ctx.eval_mult_keygen(keys)
    ctx.bootstrap_setup(budget, slots)
    ctx.bootstrap_keygen(keys, slots)
    return ctx, keys


@pytest.fixture(scope="module")
def env():
    return make(12, 26, 8, (2, 2))


def deep(ctx, keys, v, slots, levels):
    """an encryption of v pushed down to `levels` consumed levels by multiplications with 1.0"""
    c = ctx.encrypt(keys, ctx.make_plaintext(v, 1, 0, slots))
    while c.level < levels:
        c = ctx.mult(c, 1.0)
    return c


def test_bootstrap_steps(env):
    ctx, keys = env
    s, K = 8, 512
    z = np.array([0.5, -0.25, 0.125, 0.9, -0.7, 0.0, 0.33, -0.6])
    c = deep(ctx, keys, z, s, ctx.depth - 1)
    assert ctx.bootstrap_depth(s) == 1 + 2 + 7 + 6 + 2
    # step 1: one limb, message intact under the small scale q0 2^-c
    c1 = ctx.bootstrap(c, 1)
    i1 = c1.info()
    assert i1["limbs"] == 1 and abs(i1["scale"] / (ctx.moduli[0] * 2.0 ** -CORRECTION) - 1) < 1e-9
    assert np.abs(ctx.decrypt(keys, c1, s) - z).max() < 1e-6
    # step 3: slots = t / (q0 K) with t = msg_scale m + q0 I: the fractional part of K x carries the coefficients
    w = np.linalg.solve(special_matrix(s), z.astype(complex))
    br = bitrev(s)
    coef = np.concatenate([w.real[br], w.imag[br]])
    c3 = ctx.bootstrap(c, 3)
    assert c3.info()["slots"] == 2 * s
    x = ctx.decrypt(keys, c3, 2 * s)
    assert np.abs(x).max() < 1.0
    x = x * (1 + 1.0 / (4 * K))                     # the plan folds 1 / (1 + delta) into CoeffsToSlots
    frac = K * x - np.round(K * x)
    assert np.abs(frac * 2.0 ** CORRECTION - coef).max() < 1e-4
    # step 4: sin(2 pi t / q0) ~ 2 pi 2^-c m
    y = ctx.decrypt(keys, ctx.bootstrap(c, 4), 2 * s)
    assert np.abs(y / (2 * np.pi * 2.0 ** -CORRECTION) - coef).max() < 1e-4
    # whole thing
    out = ctx.bootstrap(c)
    io = out.info()
    assert io["slots"] == s and io["level"] + io["nsd"] - 1 == ctx.bootstrap_depth(s)
    err = np.abs(ctx.decrypt(keys, out, s) - z).max()
    print("bootstrap max error 2^%.1f, output level %d" % (np.log2(err), io["level"]))
    assert err < 2.0 ** -16


def test_modraise_is_the_centred_lift_bit_for_bit(env):
    """ModRaise against its integer definition: inverse NTT of the one-limb ciphertext, centred representative in
    (-q0/2, q0/2], residues modulo every limb of level 1, forward NTT (the engine's own NTT is bit-exact vs the oracle,
    tests/test_gpu_primitives.py)."""
    ctx, keys = env
    c = deep(ctx, keys, np.linspace(-0.8, 0.8, 8), 8, ctx.depth - 1)
    small = ctx.ct_export(ctx.bootstrap(c, 1))
    raised = ctx.ct_export(ctx.bootstrap(c, 5))
    q0 = ctx.moduli[0]
    assert raised[0].shape[0] == ctx.L - 1
    for poly in range(2):
        coef = [int(v) for v in ctx.raw_ntt(0, small[poly][0], inverse=True)]
        cent = [v - q0 if v > q0 // 2 else v for v in coef]
        for limb in (0, 1, 7, ctx.L - 2):
            q = ctx.moduli[limb]
            want = ctx.raw_ntt(limb, np.array([v % q for v in cent], dtype=np.uint64))
            assert np.array_equal(raised[poly][limb], want), (poly, limb)


def test_double_hoisted_stages_match_single_hoisting():
    """FHEB200_BOOT_HOIST=single evaluates every rotation of a linear-transform stage with its own ModDown; the default
    sums the stage in the extended basis and pays for one.  Both must decrypt to the same values."""
    import os
    import subprocess
    import sys
    code = ("import numpy as np, sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
            "from test_gpu_bootstrap import make, deep\n"
            "ctx, keys = make(13, 26, 64, (3, 3))\n"
            "z = np.random.default_rng(2).uniform(0, 1, 64)\n"
            "out = ctx.bootstrap(deep(ctx, keys, z, 64, ctx.depth - 1))\n"
            "print(' '.join('%%.9f' %% v for v in ctx.decrypt(keys, out, 64)))\n") % (
                os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    res = {}
    for mode in ("double", "single"):
        env2 = dict(os.environ, FHEB200_BOOT_HOIST=mode)
        o = subprocess.run([sys.executable, "-c", code], env=env2, capture_output=True, text=True, timeout=300)
        assert o.returncode == 0, o.stderr[-2000:]
        res[mode] = np.array([float(v) for v in o.stdout.split()])
    z = np.random.default_rng(2).uniform(0, 1, 64)
    assert np.abs(res["double"] - z).max() < 2.0 ** -16 and np.abs(res["single"] - z).max() < 2.0 ** -16
    assert np.abs(res["double"] - res["single"]).max() < 2.0 ** -16


def test_bootstrap_then_compute(env):
    ctx, keys = env
    s = 8
    z = np.linspace(0.05, 0.95, s)
    c = deep(ctx, keys, z, s, ctx.depth - 1)
    b = ctx.bootstrap(c)
    sq = ctx.mult(ctx.mult(b, b), b)
    assert np.abs(ctx.decrypt(keys, sq, s) - z ** 3).max() < 1e-4
    # a second bootstrapping of a bootstrapped value
    d = b
    while d.level < ctx.depth - 1:
        d = ctx.mult(d, 1.0)
    b2 = ctx.bootstrap(d)
    assert np.abs(ctx.decrypt(keys, b2, s) - z).max() < 1e-4


def test_bootstrap_refuses_without_spare_level(env):
    import fhe_sorting_b200 as F
    ctx, keys = env
    c = deep(ctx, keys, np.zeros(8), 8, ctx.depth)   # last level with a pending rescale: nothing left to scale down with
    assert c.info()["nsd"] == 2 and c.info()["limbs"] == 2
    with pytest.raises(F.FheError):
        ctx.bootstrap(c)


@pytest.mark.parametrize("logn,slots,budget", [(13, 64, (3, 3)), (14, 1024, (5, 5))])
def test_bootstrap_sizes(logn, slots, budget):
    ctx, keys = make(logn, 1 + budget[0] + 13 + budget[1] + 3, slots, budget, seed=9)
    z = np.random.default_rng(slots).uniform(0, 1, slots)
    c = deep(ctx, keys, z, slots, ctx.depth - 1)
    out = ctx.bootstrap(c)
    err = np.abs(ctx.decrypt(keys, out, slots) - z).max()
    print("slots %d: bootstrap max error 2^%.1f" % (slots, np.log2(err)))
    assert err < 2.0 ** -13


def test_meta_bts_second_iteration_removes_the_first_pass_error(env):
    """EvalBootstrap(ct, 2, precision) (reference call: src/sort_algo.h:1110-1111 EvalBootstrap(result, 2, 20)): the second
    pass bootstraps the scaled error of the first, so the output error drops by orders of magnitude; levels: one more than
    the single pass (the 2^-p constant)."""
    ctx, keys = env
    z = np.array([0.5, -0.25, 0.125, 0.9, -0.7, 0.0, 0.33, -0.6])
    c = deep(ctx, keys, z, 8, ctx.depth - 2)
    one = ctx.bootstrap(c)
    two = ctx.bootstrap(c, iterations=2, precision=20)
    e1 = np.abs(ctx.decrypt(keys, one, 8) - z).max()
    e2 = np.abs(ctx.decrypt(keys, two, 8) - z).max()
    print("bootstrap error: one pass 2^%.1f, Meta-BTS 2^%.1f" % (np.log2(e1), np.log2(e2)))
    assert e1 < 2.0 ** -12
    assert e2 < e1 / 64
    assert two.level <= one.level + 2
