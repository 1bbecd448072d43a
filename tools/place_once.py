#!/usr/bin/env python
"""rotationIndexCheckN of DirectSort<N> on plain ranks (the Paterson-Stockmeyer heavy phase), twice; prints the launch
count at the start of the second run so that ncu can skip to it: python tools/place_once.py N"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input  # noqa: E402

n = int(sys.argv[1])
s = DirectSortSession(n, 17, seed=1000 + n)
x = uniform_gap_input(n)
ranks = np.argsort(np.argsort(x)).astype(float)
s.place(x, ranks)
print("second run starts at launch: %d" % F.launch_count(), file=sys.stderr, flush=True)
t0 = time.perf_counter()
out, ms, level = s.place(x, ranks)
print("N=%d place %.1f ms err %.2e launches %d" % (n, 1e3 * (time.perf_counter() - t0), np.abs(out - np.sort(x)).max(),
                                                   F.launch_count()), flush=True)
