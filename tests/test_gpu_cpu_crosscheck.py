"""GPU engine against an independent CPU run of the whole sort.  Same drivers, same key seed, same input; on one side
libfheb200.so (CUDA), on the other oracle/libfheb200_cpu.so (the oracle's primitives under a CPU scheme layer).  Encryption
randomness differs (the CPU layer draws its own), so the comparison is on the decrypted values: both within the reference's
0.01 of the truth (tests/DirectSortTest.cpp:169) and within 1e-5 of EACH OTHER - approximation error of the sign polynomials
is common to both, what is left is CKKS noise."""
import numpy as np
import pytest

import refdrivers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,logring", [(8, 12), (16, 12), (32, 13)])
def test_direct_sort_gpu_equals_cpu(n, logring, capfd):
    from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
    x = uniform_gap_input(n)
    g = DirectSortSession(n, logring, seed=900 + n)
    gout, _, glevel = g.sort(x)
    grank, _, _ = g.rank(x)
    c = DirectSortSession(n, logring, seed=900 + n, lib=refdrivers.cpu_library("own"))
    cout, _, clevel = c.sort(x)
    crank, _, _ = c.rank(x)
    assert (g.mult_depth, g.num_rotation_keys) == (c.mult_depth, c.num_rotation_keys)
    assert glevel == clevel == g.mult_depth
    assert np.abs(gout - np.sort(x)).max() < 0.01 and np.abs(cout - np.sort(x)).max() < 0.01
    assert np.abs(gout - cout).max() < 1e-5, "sorted values: GPU vs CPU"
    assert np.abs(grank - crank).max() < 1e-4, "ranks: GPU vs CPU"
    g.close(), c.close()


def test_reference_drivers_gpu_equals_cpu(capfd):
    """The reference's unmodified DirectSort on both backends."""
    from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
    n = 16
    x = uniform_gap_input(n, seed=5)
    g = DirectSortSession(n, 12, seed=31, lib=refdrivers.library())
    c = DirectSortSession(n, 12, seed=31, lib=refdrivers.cpu_library("reference"))
    gout, _, glevel = g.sort(x)
    cout, _, clevel = c.sort(x)
    assert glevel == clevel == g.mult_depth == c.mult_depth
    assert np.abs(gout - cout).max() < 1e-5
    assert np.abs(gout - np.sort(x)).max() < 0.01
    g.close(), c.close()


def test_mehp24_gpu_equals_cpu(capfd):
    from fhe_sorting_b200.sort import Mehp24SortSession, uniform_gap_input
    x = uniform_gap_input(16)
    g = Mehp24SortSession(16, 12, seed=44)
    c = Mehp24SortSession(16, 12, seed=44, lib=refdrivers.cpu_library("own"))
    gout, _, glevel = g.sort(x)
    cout, _, clevel = c.sort(x)
    assert glevel == clevel and g.mult_depth == c.mult_depth
    assert np.abs(gout - cout).max() < 1e-5
    assert np.abs(gout - np.sort(x)).max() < 0.01
    g.close(), c.close()
