"""Host-side logic of the sort drivers that needs no GPU: rotation-key schedule, index decomposition, Chebyshev tables."""
import itertools
import os

import numpy as np
import pytest

from fhe_sorting_b200 import sort as S


def test_rotation_indices_cover_the_schedule():
    for n in (4, 16, 64, 256, 1024):
        idx = S.rotation_indices(n, 17)
        P = min(n, 65536 // n)
        assert all(0 < r < n * P for r in idx) and idx == sorted(set(idx))
        # folds over the P copies
        span = n * P // 2
        while span >= n:
            assert span in idx
            span //= 2
    assert len(S.rotation_indices(1024, 17)) == 161   # same key count as the reference's N = 1024 table


@pytest.mark.skipif(not os.path.exists("/root/reference/src/sort_algo.h"), reason="reference checkout not present")
def test_rotation_indices_match_reference_tables_up_to_128():
    import re
    src = open("/root/reference/src/sort_algo.h").read()
    body = src[src.index("static void getSizeParameters"):src.index("parameters.SetScalingModSize(modSize)")]
    for m in re.finditer(r"case (\d+):(.*?)break;", body, re.S):
        n = int(m.group(1))
        if n > 128:
            continue
        ref = {int(x) for x in re.findall(r"\d+", re.search(r"rotations = \{(.*?)\};", m.group(2), re.S).group(1))}
        slots = n * min(n, 65536 // n)
        assert {r for r in ref if r % slots} == set(S.rotation_indices(n, 17))


def test_decompose_sums_to_rotation():
    """tests/DecomposeTest.cpp:64-74: the signed steps always add up to the requested rotation (modulo N for the NAF form,
    whose -N/2 digit is emitted as +N/2, src/rotation.h:124-128; Decomposer<1024> here)."""
    keys = [1, 2, 4, 8, 16, 32, 64, 512]
    for algo, rot in itertools.product(("NAF", "BNAF", "BINARY"), list(range(1, 130)) + [511, 513, 1000]):
        steps = S.decompose(keys, rot, 1 << 20, algo)
        assert (sum(steps) - rot) % 1024 == 0, (algo, rot, steps)
    assert S.decompose(keys, 5, 1 << 20, "BINARY") == [4, 1]          # largest step first
    assert S.decompose(keys, 7, 1 << 20, "NAF") == [8, -1]
    assert S.decompose(keys, 7, 1 << 20, "BNAF") == [8, -1]
    assert S.decompose(keys, 3, 1 << 20, "BNAF") == [4, -1]           # balanced recoding: digit 1 followed by digit 1 borrows
    assert S.decompose(keys, 5, 1 << 20, "BNAF") == [4, 1]
    assert S.decompose(keys, 600, 1 << 20, "NAF")[0] == 512           # the largest key is peeled off first
    assert S.decompose(keys, 96, 64, "BINARY") == [32]                 # steps that are multiples of the wrap length are dropped


def test_decompose_equals_the_reference_decomposer():
    """Every rotation 1..1100, the three digit forms, two key lists: same steps in the same order as the reference's own
    Decomposer<1024> (src/rotation.h:29-165, compiled unmodified into oracle/_ref/libref_sort.so)."""
    import refdrivers
    ref = refdrivers.library()
    for keys in ([1, 2, 4, 8, 16, 32, 64, 512], [-1, -2, -4, -8, 1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048]):
        for algo in ("NAF", "BNAF", "BINARY"):
            for rot in range(1, 1101):
                for wrap in (128, 1 << 20):
                    assert S.decompose(keys, rot, wrap, algo) == S.decompose(keys, rot, wrap, algo, lib=ref), (keys, algo, rot, wrap)


def test_doubled_sinc_table_is_a_one_hot_at_rank_offsets():
    n = 16
    c = S.doubled_sinc_coefficients(n)
    assert 100 < len(c) < 160
    assert all(float("%g" % v) == v for v in c)          # 6 significant digits, as the reference's generated header
    k = np.arange(-2 * n + 1, n)                        # index - rank - r ranges over (-2N, N)
    x = k / (2.0 * n)
    y = np.polynomial.chebyshev.chebval(x, np.concatenate([[c[0] / 2], c[1:]]))
    want = ((k == 0) | (k == -n)).astype(float)
    assert np.abs(y - want).max() < 2e-3


def test_tables_equal_the_reference_generator_output():
    """oracle/_ref/gen/*.h are written by the reference's own utils/generate_cheb_*_coeffs.cpp (compiled unmodified
    against the shim); the tables built at run time must be identical, value for value."""
    import re
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "gen",
                        "generated_doubled_sinc_coeffs.h")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    txt = open(path).read()
    for n in (4, 16, 64, 256):
        m = re.search(r"generatedDoubledSincCoefficients_%d = \{(.*?)\};" % n, txt, re.S)
        ref = np.array([float(v) for v in m.group(1).replace("\n", " ").split(",")])
        assert np.array_equal(ref, S.doubled_sinc_coefficients(n))


def test_tables_equal_golden_reference_generator_output():
    """tests/golden/chebyshev_tables.json: tables printed by the reference's own generators (tools/make_golden.py)."""
    import json
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "chebyshev_tables.json")))
    for n, want in gold["doubled_sinc"].items():
        assert np.array_equal(np.array(want), S.doubled_sinc_coefficients(int(n)))
    for n, want in gold["sinc"].items():
        assert np.array_equal(np.array(want), S.sinc_coefficients(int(n)))
