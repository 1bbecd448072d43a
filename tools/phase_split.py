#!/usr/bin/env python
"""DirectSort N: time of the rank phase (constructRank) against the whole sort, with operation counters: python tools/phase_split.py N"""
import sys
import numpy as np
sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
from fhe_sorting_b200.lib import op_counts
n = int(sys.argv[1])
s = DirectSortSession(n, 17, seed=1000 + n)
x = uniform_gap_input(n)
s.sort(x)
for name, fn in (("rank", s.rank), ("sort", s.sort)):
    c0 = op_counts()
    out, ms, level = fn(x)
    c1 = op_counts()
    print("N=%d %s: %.1f ms, ops (rescale, relin, merged, rotations) %s" % (n, name, ms, tuple(b - a for a, b in zip(c0, c1))), flush=True)
