#!/usr/bin/env python
"""One encrypted DirectSort at ring 2^17 (for profiling under ncu): python tools/sort_once.py N [reps]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input  # noqa: E402

n = int(sys.argv[1])
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
s = DirectSortSession(n, 17, seed=1000 + n)
x = uniform_gap_input(n)
import torch  # noqa: E402
for rep in range(reps):
    if rep == reps - 1:
        torch.cuda.profiler.start()  # ncu --profile-from-start off: the last sort only
    from fhe_sorting_b200.lib import op_counts
    c0 = op_counts()
    t0 = time.perf_counter()
    out, ms, level = s.sort(x)
    c1 = op_counts()
    print("ops (rescale, relin, merged, rotations):", tuple(b - a for a, b in zip(c0, c1)), flush=True)
    print("N=%d sort %.1f ms wall %.1f ms err %.2e" % (n, ms, 1e3 * (time.perf_counter() - t0), np.abs(out - np.sort(x)).max()),
          flush=True)
