"""Host-side checks of the k-way network geometry (fhe_sorting_b200/host/kway.h) against golden vectors written by the
reference's own src/k-way/Masking.cpp (tools/make_golden.py -> tests/golden/kway_stages.json) and the reference's
known answers (tests/k-way/MaskingTest.cpp:45-105).  No GPU."""
import json
import os

import numpy as np
import pytest

from fhe_sorting_b200 import sort as S

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kway_stages.json")))


@pytest.mark.parametrize("key", sorted(GOLD))
def test_stage_geometry_equals_reference(key):
    k, m = (int(x) for x in key.split("^"))
    assert S.kway_num_stages(k, m) == len(GOLD[key])
    for st, want in enumerate(GOLD[key]):
        geo, idx = S.kway_stage(k, m, st, k ** m)
        assert list(geo) == want["geometry"], (key, st)
        assert np.array_equal(idx, np.array(want["indices"])), (key, st)


def test_reference_known_answers():
    # MaskingTest.cpp:45-58 sortType(5, 3, 0) == (0, 0, 0) ; :97-105 getRotateDistance
    assert S.kway_stage(5, 3, 0, 125)[0][:3] == (0, 0, 0)
    geo = S.kway_stage(5, 3, 5, 125)[0]
    assert geo[0] >= 0 and geo[1] >= 0 and geo[2] <= 5 // 2 + 1
    # getRotateDistance(2,1,0) = 2 ; (3,1,1) = 6 ; (5,1,3) = 5 : read off stages that have those (k, logDist, slope)
    seen = {}
    for k, m in ((2, 3), (3, 3), (5, 3)):
        for st in range(S.kway_num_stages(k, m)):
            g = S.kway_stage(k, m, st, k ** m)[0]
            seen[(k, g[1], g[2])] = g[3]
    assert seen[(2, 1, 0)] == 2 and seen[(3, 1, 1)] == 6 and seen[(5, 1, 3)] == 5


def test_network_sorts_in_the_clear():
    """The stage geometry drives a plain-numpy run of the same data flow as kwaySort::Sorter (k = 2 and 3): neighbour
    comparison bits from one rotate-and-compare, small sorters on slot-aligned copies, masked reassembly."""
    for k, m in ((2, 4), (3, 3)):
        n = k ** m
        slots = 1 << (n - 1).bit_length()
        rng = np.random.default_rng(n)
        x = np.zeros(slots)
        x[:n] = rng.permutation(n) / n
        for st in range(S.kway_num_stages(k, m)):
            (mm, logd, slope, shift), idx = S.kway_stage(k, m, st, n)
            size = np.zeros(slots, dtype=int)
            pos = np.zeros(slots, dtype=int)
            size[:n], pos[:n] = idx[0], idx[1]
            inner = (pos < size).astype(float)
            rot = np.roll(x * inner, shift)
            for s in range(1, k + 1):
                last = ((size == s) & (pos == s)).astype(float)
                rot = rot + np.roll(x * last, -(s - 1) * shift)
            bits = (x > rot).astype(float)
            rest = x * ((size == 0) | ((size == 1) & (pos == 1)))
            out = rest.copy()
            for s in range(2, k + 1):
                first = ((size == s) & (pos == 1)).astype(float)
                if not first.any():
                    continue
                members = np.stack([np.roll(x, -p * shift) for p in range(s)])
                srt = np.sort(members, axis=0)
                for p in range(s):
                    out = out + np.roll(srt[p] * first, p * shift)
            x = out
        assert np.array_equal(x[:n], np.sort(x[:n])) and np.allclose(np.sort(x[:n]), np.arange(n) / n)
