"""MEHP24 ranking sort through the C++ host drivers (host/mehp24.h over the lbcrypto shim): the assertion of the
reference's tests/mehp24/Mehp24SortTest.cpp:184 (max error < 0.01) with its per-N depth and sign configuration, -m gpu."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,logn", [(4, 12), (8, 12), (16, 13), (32, 13), (64, 13)])
def test_mehp24_sortfg_small_rings(n, logn):
    from fhe_sorting_b200.sort import Mehp24SortSession, uniform_gap_input
    s = Mehp24SortSession(n, logn, seed=300 + n)
    x = uniform_gap_input(n, seed=n)
    out, ms, level = s.sort(x)
    assert np.abs(out - np.sort(x)).max() < 0.01
    assert level <= s.mult_depth
    s.close()


def test_mehp24_n256_reference_parameters():
    """config 4 of BASELINE.json: N = 256, ring 2^17, 65536 slots, depth 49, CompositeSign(3, 4, 2), dg_i = 4, df_i = 2."""
    from fhe_sorting_b200.sort import Mehp24SortSession, uniform_gap_input
    s = Mehp24SortSession(256, 17, seed=9)
    assert s.mult_depth == 49
    x = uniform_gap_input(256)
    out, ms, level = s.sort(x)
    assert np.abs(out - np.sort(x)).max() < 0.01
    assert level <= 49
    s.close()


def test_mehp24_reference_drivers_agree():
    """The reference's own src/mehp24/*.cpp compiled unmodified against the shim (oracle/_ref) and run on the engine."""
    import refdrivers
    from fhe_sorting_b200.sort import Mehp24SortSession, uniform_gap_input
    if not refdrivers.available():
        pytest.skip("oracle/_ref/libref_sort.so not built (no reference checkout on the build machine)")
    n, logn = 16, 13
    x = uniform_gap_input(n, seed=4)
    ref = Mehp24SortSession(n, logn, seed=5, lib=refdrivers.library())
    out_ref, _, level_ref = ref.sort(x)
    ref.close()
    ours = Mehp24SortSession(n, logn, seed=5)
    out, _, level = ours.sort(x)
    ours.close()
    assert np.abs(out_ref - np.sort(x)).max() < 0.01
    assert level == level_ref
    assert np.abs(out - out_ref).max() < 1e-3
