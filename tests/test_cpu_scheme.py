"""The CPU scheme layer (oracle/cpu_scheme.cpp -> oracle/libfheb200_cpu.so): the C ABI of include/fhe_b200.h restated on the
host cores over the oracle's primitives - FLEXIBLEAUTOEXT level/scale bookkeeping, EvalMult + relinearise, rescale,
EvalRotate, hoisted rotations, Paterson-Stockmeyer Chebyshev - so that the sort drivers (this repository's and, where
/root/reference was present at build time, the reference's unmodified ones) run a whole encrypted DirectSort with no CUDA.
These tests pin that layer on the CPU (DirectSortTest.cpp's assertions); tests/test_gpu_cpu_crosscheck.py then holds the
GPU engine against it."""
import ctypes as C
import os

import numpy as np
import pytest

import refdrivers
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input

HERE = os.path.dirname(os.path.abspath(__file__))
CPU_LIB = os.path.join(os.path.dirname(HERE), "oracle", "libfheb200_cpu.so")


def test_cpu_library_exports_the_scheme_abi():
    L = C.CDLL(CPU_LIB)
    for name in ("fheb200_ctx_create", "fheb200_keygen", "fheb200_eval_mult_keygen", "fheb200_eval_rotate_keygen",
                 "fheb200_make_packed_plaintext", "fheb200_encrypt", "fheb200_decrypt", "fheb200_eval_add", "fheb200_eval_sub",
                 "fheb200_eval_mult", "fheb200_eval_mult_plain", "fheb200_eval_mult_const", "fheb200_eval_rotate",
                 "fheb200_fast_rotation_precompute", "fheb200_eval_fast_rotation", "fheb200_eval_chebyshev_series",
                 "fheb200_eval_poly", "fheb200_eval_rescale", "fheb200_ct_info"):
        assert hasattr(L, name), name
    L.fheb200_version.restype = C.c_char_p
    assert b"cpu" in L.fheb200_version()


@pytest.mark.parametrize("which", ["own", "reference"])
def test_direct_sort_on_the_host_cores(which, capfd):
    """tests/DirectSortTest.cpp:128,169 with no GPU: sorted to 0.01, final level == multiplicative depth."""
    L = refdrivers.cpu_library(which)
    n = 8
    s = DirectSortSession(n, 12, seed=77, lib=L)
    x = uniform_gap_input(n)
    out, _, level = s.sort(x)
    assert np.abs(out - np.sort(x)).max() < 1e-4
    assert level == s.mult_depth
    ranks, _, _ = s.rank(x)
    assert np.abs(ranks - np.argsort(np.argsort(x))).max() < 1e-3
    s.close()


def test_cpu_scheme_rejects_unseeded_keys():
    L = refdrivers.cpu_library("own")
    with pytest.raises(Exception):
        DirectSortSession(8, 12, seed=0, lib=L)     # the oracle has no entropy source: seed 0 (OS entropy) is refused, loudly


def test_mehp24_on_the_host_cores(capfd):
    """tests/mehp24/Mehp24SortTest.cpp's assertions with no GPU (EvalPolyLinear + the g/f composite sign + rankings)."""
    from fhe_sorting_b200.sort import Mehp24SortSession
    s = Mehp24SortSession(8, 12, seed=7, lib=refdrivers.cpu_library("own"))
    x = uniform_gap_input(8)
    out, _, level = s.sort(x)
    assert np.abs(out - np.sort(x)).max() < 1e-4
    assert level <= s.mult_depth
    s.close()
