"""Known-answer vectors of the reference's primitive-level tests, through the C++ host classes of this repository and,
where oracle/_ref was built, through the reference's own unmodified classes on the same engine (-m gpu):

  tests/RotationTest.cpp:65-130      RotationComposer (NAF) and RotationTree<8>
  tests/k-way/EvalUtilsTest.cpp:62-158, SortUtilsTest.cpp:68-113, SorterTest.cpp:67-260
  src/sign.cpp:610-651               the sign() dispatcher: Tanh, NaiveDiscrete, SignumPolycircuit (utils/sign_characterizer.cpp)
"""
import numpy as np
import pytest

import refdrivers
from fhe_sorting_b200 import sort as S

pytestmark = pytest.mark.gpu

ROT_KEYS = [-1, -2, -4, -8, -16, -32, 1, 2, 4, 8, 16, 32, 64, 512]        # tests/RotationTest.cpp:46


def _libs():
    return [("ours", None)] + ([("reference", refdrivers.library())] if refdrivers.available() else [])


@pytest.mark.parametrize("who,lib", _libs(), ids=lambda v: v if isinstance(v, str) else "")
def test_rotation_tree_table(who, lib):
    """tests/RotationTest.cpp:77-109: RotationTree<8>, NAF, buildTree(-4, 4), every rotation against the table (1e-6).
    Batch size 8 here: the reference fixture packs the 8 values into 128 slots, where a rotation brings in the zero padding
    and the table (a cyclic shift of 8 values) cannot hold."""
    s = S.ToolkitSession(13, 4, 59, 8, ROT_KEYS, seed=7, lib=lib)
    x = np.arange(1.0, 9.0)
    # the reference's buildTree(start, end) decomposes with wrap length `end` = 4 (src/rotation.h:273), which drops the +-4
    # step from every path, while treeRotate decomposes with the slot count: for |r| >= 3 it finds no node for the 4-step,
    # prints "Child node not found" and returns the input unrotated.  Compared with the reference's class on -2..2 only.
    rots = list(range(-4, 5)) if who == "ours" else list(range(-2, 3))
    got = s.tree_rotate(8, x, -4, 4, rots, "NAF")
    for r, row in zip(rots, got):
        assert np.abs(row - np.roll(x, -r)).max() < 1e-6, "rotation %d (%s drivers)" % (r, who)
    s.close()


def test_rotation_tree_bigger_and_other_digit_forms():
    s = S.ToolkitSession(13, 4, 59, 128, ROT_KEYS, seed=8)
    x = np.random.default_rng(1).uniform(-1, 1, 128)
    rots = [-31, -17, -1, 1, 3, 7, 15, 21, 31]
    for algo in ("NAF", "BNAF", "BINARY"):
        got = s.tree_rotate(128, x, -31, 31, rots, algo)
        for r, row in zip(rots, got):
            assert np.abs(row - np.roll(x, -r)).max() < 1e-6, (algo, r)
    s.close()


@pytest.mark.parametrize("who,lib", _libs(), ids=lambda v: v if isinstance(v, str) else "")
def test_composer_forward_and_backward(who, lib):
    """tests/RotationTest.cpp:111-130 RotateForwardAndBackward: rotate by r then by -r, r in -128..128, 1e-5."""
    s = S.ToolkitSession(13, 4, 59, 128, ROT_KEYS, seed=9, lib=lib)
    x = np.random.default_rng(2).permutation(25500)[:128] * 0.01
    for r in (-128, -127, -65, -33, -3, -1, 0, 1, 2, 5, 31, 63, 64, 65, 96, 127, 128):
        fwd = s.composed_rotate(x, r)
        assert np.abs(fwd - np.roll(x, -r)).max() < 1e-5, (who, r)
        assert np.abs(s.composed_rotate(fwd, -r) - x).max() < 1e-5, (who, r)
    s.close()


@pytest.fixture(scope="module")
def kway_fixture():
    """tests/k-way/SorterTest.cpp:12-47: ring 2^12, depth 59, scaling 59, batch 16, keys +-1..15, bootstrapping {4,4}"""
    keys = [v for i in range(1, 16) for v in (i, -i)]
    s = S.ToolkitSession(12, 59, 59, 16, keys, bootstrap_budget=4, seed=11)
    yield s
    s.close()


def _pad(v, fill=0.0, n=16):
    out = np.full(n, fill)
    out[:len(v)] = v
    return out


def test_kway_eval_utils(kway_fixture):
    s = kway_fixture
    x = np.array([1.0, 2.0, 3.0, 4.0])
    assert np.abs(s.kway_unit(1, 2, 2, _pad(x), param=3)[0][:4] - 3 * x).max() < 0.1       # EvalUtilsTest.cpp:62-74
    assert np.abs(s.kway_unit(1, 2, 2, _pad(x), param=-2)[0][:4] + 2 * x).max() < 0.1      # :76-88
    y = np.array([2.0, 3.0, 4.0, 5.0])
    assert np.abs(s.kway_unit(2, 2, 2, _pad(x), _pad(y))[0][:4] - x * y).max() < 0.1       # :90-108
    assert np.abs(s.kway_unit(3, 2, 2, _pad(x))[0][:4] - x * x).max() < 0.1
    v = np.arange(1.0, 17.0)
    assert np.abs(s.kway_unit(4, 2, 2, v, param=3)[0] - np.roll(v, -3)).max() < 0.1        # :110-130
    assert np.abs(s.kway_unit(5, 2, 2, v, param=2)[0] - np.roll(v, 2)).max() < 0.1
    # multByInt costs no level
    _, before, after = s.kway_unit(1, 2, 2, _pad(x), param=5)
    assert before == after


def test_kway_level_guard_bootstraps_only_when_needed(kway_fixture):
    """EvalUtilsTest.cpp:132-158: a ciphertext encoded at level 40 of depth 59 keeps its level when 2 levels are asked for
    and is bootstrapped (level drops) when more are asked for than remain; values survive within 0.2."""
    s = kway_fixture
    x = _pad([0.1, 0.2, 0.3, 0.4])
    out, before, after = s.kway_unit(6, 2, 2, x, param=2, aux=40)
    assert before == after == 40 and np.abs(out[:4] - x[:4]).max() < 0.2
    out, before, after = s.kway_unit(6, 2, 2, x, param=20, aux=40)
    assert after < before and np.abs(out[:4] - x[:4]).max() < 0.2


def test_kway_sort_utils(kway_fixture):
    s = kway_fixture
    a, b, c = [2.0, 4.0, 6.0, 8.0], [1.0, 5.0, 3.0, 7.0], [1.0, 0.0, 1.0, 1.0]
    out = s.kway_unit(10, 2, 2, _pad(a), _pad(b), _pad(c))[0]
    assert np.abs(out[:4] - [2.0, 5.0, 6.0, 8.0]).max() < 0.1                              # SortUtilsTest.cpp:68-89
    a, b, c = [5.0, 2.0, 8.0, 1.0], [3.0, 6.0, 4.0, 7.0], [1.0, 0.0, 1.0, 0.0]
    out = s.kway_unit(11, 2, 2, _pad(a), _pad(b), _pad(c))[0]
    assert np.abs(out[:4] - [3.0, 2.0, 4.0, 1.0]).max() < 0.1                              # :91-113 (min, then max)
    assert np.abs(out[16:20] - [5.0, 6.0, 8.0, 7.0]).max() < 0.1


def test_kway_run_sorters(kway_fixture):
    s = kway_fixture
    # SorterTest.cpp:67-98 RunTwoSorter
    x = [5.0, 2.0, 8.0, 1.0, 3.0, 6.0, 4.0, 7.0]
    comp = [1.0, 0.0, 1.0, 0.0, 0.0, 1.0, 0.0, 1.0]
    idx = np.zeros((2, 16), dtype=np.int32)
    idx[0, 0:8:2], idx[1, 0:8:2] = 2, 1
    out = s.kway_unit(22, 2, 2, _pad(x), _pad(comp), indices=idx, param=1)[0]
    assert np.abs(out[:8] - [2.0, 5.0, 1.0, 8.0, 3.0, 6.0, 4.0, 7.0]).max() < 0.1
    # :100-161 RunThreeSorter (the HEAAN example)
    x = [0.840188, 0.394383, 0.783099, 0.79844, 0.911647, 0.197551, 0.335223, 0.76823, 0.277775]
    comp = _pad([1.0, 0.0, 1.0, 1.0, 1.0, 0.0, 1.0, 1.0, 0.0], 0.5)
    idx = np.zeros((2, 16), dtype=np.int32)
    idx[0, :9], idx[1, :9] = 3, np.arange(9) % 3 + 1
    out = s.kway_unit(23, 3, 2, _pad(x), comp, indices=idx, param=1)[0]
    want = [0.394404, 0.783115, 0.840147, 0.197544, 0.79844, 0.911647, 0.277775, 0.335223, 0.76823]
    assert np.abs(out[:9] - want).max() < 0.1
    # :163-211 RunFourSorter
    idx = np.zeros((2, 16), dtype=np.int32)
    idx[0, :4], idx[1, :4] = 4, [1, 2, 3, 4]
    out = s.kway_unit(24, 4, 2, _pad([5.0, 2.0, 8.0, 1.0]), _pad([1.0, 0.0, 1.0, 0.0], 0.5), _pad([0.0, 1.0, 1.0, 0.0], 0.5),
                      indices=idx, param=1)[0]
    assert np.abs(out[:4] - [1.0, 2.0, 5.0, 8.0]).max() < 0.1
    # :213-260 RunFiveSorter
    idx = np.zeros((2, 16), dtype=np.int32)
    idx[0, :5], idx[1, :5] = 5, [1, 2, 3, 4, 5]
    out = s.kway_unit(25, 5, 2, _pad([5.0, 2.0, 8.0, 1.0, 3.0]), _pad([1.0, 0.0, 1.0, 0.0, 1.0], 0.5),
                      _pad([1.0, 0.0, 1.0, 0.0, 0.0], 0.5), indices=idx, param=1)[0]
    assert np.abs(out[:5] - [1.0, 2.0, 3.0, 5.0, 8.0]).max() < 0.1


@pytest.fixture(scope="module")
def sign_fixture():
    """utils/sign_characterizer.cpp:26-38: depth 40, scaling 50, batch 8"""
    s = S.ToolkitSession(12, 40, 50, 8, seed=13)
    yield s
    s.close()


def test_sign_dispatcher_variants(sign_fixture):
    """Tanh (degree 1006), NaiveDiscrete (degree 119, a 0/1 step) and SignumPolycircuit (degree 1023) against the exact
    function at |x| >= 0.05 with the 0.1 tolerance of tests/SincTest.cpp / SignTest.cpp:69; depths = ceil(log2(deg + 1))."""
    s = sign_fixture
    x = np.array([-1.0, -0.6, -0.2, -0.05, 0.05, 0.3, 0.75, 1.0])
    out, level, _ = s.sign("Tanh", x)
    assert np.abs(out - np.sign(x)).max() < 0.1 and level == 10
    out, level, _ = s.sign("NaiveDiscrete", x)
    assert np.abs(out - (x >= 0)).max() < 0.1 and level == 7
    out, level, _ = s.sign("SignumPolycircuit", x)
    assert np.abs(out - np.sign(x)).max() < 0.1 and level == 11                            # k (2^m - 1) > 1023 needs ceil(log2 k) + m = 11
    out, level, _ = s.sign("CompositeSign", x, 4, 2, 2)
    assert np.abs(out - np.sign(x)).max() < 0.1 and level <= 20


def test_signum_polycircuit_matches_the_reference_code():
    """src/sign.cpp:188-608 compiled unmodified (T_0..T_1008 by EvalChebyshevSeries + a product tree for T_1009..T_1023)
    against this repository's single Paterson-Stockmeyer pass over the same coefficients."""
    ref = S.ToolkitSession(12, 40, 50, 8, seed=13, lib=refdrivers.library())
    ours = S.ToolkitSession(12, 40, 50, 8, seed=13)
    x = np.array([-0.9, -0.31, -0.04, -0.008, 0.006, 0.02, 0.4, 0.97])
    a, _, _ = ref.sign("SignumPolycircuit", x)
    b, _, _ = ours.sign("SignumPolycircuit", x)
    assert np.abs(a - b).max() < 1e-4
    ref.close(), ours.close()
