"""HKC+21 k-way sorting network with bootstrapping on the B200 engine (reference fixture and pass criterion:
tests/k-way/KWaySort2Test.cpp:14-58, 137-190 and KWaySort3Test.cpp — max error < 0.01 against std::sort)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from fhe_sorting_b200 import sort as S  # noqa: E402


def gap_input(n, seed):
    """tests/k-way/KWaySort2Test.cpp:137 getVectorWithMinDiff(N, 0, 1, (1 - 1e-8) / N), seeded"""
    rng = np.random.default_rng(seed)
    return rng.permutation(n).astype(np.float64) * ((1 - 1e-8) / n)


@pytest.mark.parametrize("k,m", [(2, 2), (2, 3), (2, 4), (3, 2), (3, 3), (5, 2), (5, 3)])
def test_kway_sort_small_ring(k, m):
    s = S.KWaySortSession(k, m, log_ring_dim=12, seed=3)
    assert s.mult_depth == 40
    x = gap_input(k ** m, 7)
    out, ms, level = s.sort(x)
    err = np.abs(out - np.sort(x)).max()
    print("k=%d M=%d: %.0f ms, %d guard bootstraps, level %d, max error 2^%.1f" % (k, m, ms, s.bootstraps(), level, np.log2(err)))
    assert err < 0.01
    s.close()


def test_kway_reference_drivers_unmodified_on_the_engine():
    """the reference's own src/k-way/*.cpp + kway_adapter.h (oracle/_ref/libref_sort.so) against this repository's kway.h"""
    import refdrivers
    libs = {"ours": None, "reference": refdrivers.library()}
    x = gap_input(8, 11)
    outs = {}
    for drivers in ("ours", "reference"):
        s = S.KWaySortSession(2, 3, log_ring_dim=12, seed=3, lib=libs[drivers])
        outs[drivers] = s.sort(x)[0]
        s.close()
    assert np.abs(outs["reference"] - np.sort(x)).max() < 0.01
    assert np.abs(outs["ours"] - outs["reference"]).max() < 1e-3
    # k = 5: four/five-sorters and the mixed 2..5 stage, reference's hand-written bit tables against the one-rule pairBit
    x = gap_input(25, 5)
    for drivers in ("ours", "reference"):
        s = S.KWaySortSession(5, 2, log_ring_dim=12, seed=3, lib=libs[drivers])
        outs[drivers] = s.sort(x)[0]
        s.close()
    assert np.abs(outs["reference"] - np.sort(x)).max() < 0.01
    assert np.abs(outs["ours"] - outs["reference"]).max() < 1e-3


def test_kway_full_parameters():
    """BASELINE config 5 parameter set (ring 2^17, 128-bit, depth 40, 59-bit scaling) at N = 16"""
    s = S.KWaySortSession(2, 4, log_ring_dim=17, seed=5)
    x = gap_input(16, 3)
    out, ms, level = s.sort(x)
    err = np.abs(out - np.sort(x)).max()
    print("ring 2^17 N=16: %.0f ms, max error 2^%.1f, level %d" % (ms, np.log2(err), level))
    assert err < 0.01
    s.close()


@pytest.mark.parametrize("n", [4, 8])
def test_bitonic_sort_with_bootstrapping(n):
    """tests/BitonicSortTest.cpp: ring 2^12, depth 58, 59-bit scaling, budget {3,3}, values in [0, 255], CompositeSign(4,3,3);
    pass criterion of the reference: max error < 1, no error above 0.1"""
    s = S.BitonicSortSession(n, log_ring_dim=12, seed=4)
    assert s.mult_depth == 58
    rng = np.random.default_rng(n)
    x = rng.permutation(255)[:n].astype(np.float64) + 0.25
    out, ms, level = s.sort(x)
    err = np.abs(out - np.sort(x))
    print("bitonic N=%d: %.0f ms, %d bootstraps, level %d, max error %.3g" % (n, ms, s.bootstraps(), level, err.max()))
    assert err.max() < 0.1
    assert s.bootstraps() >= 1
    s.close()
