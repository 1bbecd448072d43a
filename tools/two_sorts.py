#!/usr/bin/env python
"""Host-bound or GPU-bound?  Two DirectSort sessions (two contexts, two host threads) sorting concurrently on ONE GPU against one
alone: python tools/two_sorts.py N"""
import sys
import threading
import time
import numpy as np
sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
n = int(sys.argv[1])
x = uniform_gap_input(n)
S = [DirectSortSession(n, 17, seed=1000 + n + i) for i in range(2)]
for s in S:
    s.sort(x)
t0 = time.perf_counter(); S[0].sort(x); alone = time.perf_counter() - t0
res = [None, None]
def run(i):
    t = time.perf_counter(); out, ms, lv = S[i].sort(x); res[i] = (time.perf_counter() - t, float(np.abs(out - np.sort(x)).max()))
th = [threading.Thread(target=run, args=(i,)) for i in range(2)]
t0 = time.perf_counter()
for t in th: t.start()
for t in th: t.join()
both = time.perf_counter() - t0
print("N=%d: one sort alone %.3f s; two concurrent sorts %.3f s wall (%.3f / %.3f s each, errors %.1e %.1e) -> %.2fx the single time"
      % (n, alone, both, res[0][0], res[1][0], res[0][1], res[1][1], both / alone), flush=True)
