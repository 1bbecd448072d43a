#!/usr/bin/env python
"""One k-way sort at the reference's parameters: python tools/kway_once.py K M [log_ring_dim] [drivers]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from fhe_sorting_b200 import sort as S
k, m = int(sys.argv[1]), int(sys.argv[2])
logn = int(sys.argv[3]) if len(sys.argv) > 3 else 17
drivers = sys.argv[4] if len(sys.argv) > 4 else "ours"
t0 = time.time()
lib = None
if drivers == "reference":
    sys.path.insert(0, "tests")
    import refdrivers
    lib = refdrivers.library()
s = S.KWaySortSession(k, m, log_ring_dim=logn, seed=5, lib=lib)
print("setup %.1f s (%d rotation keys requested by the network)" % (time.time() - t0, s.num_rotation_keys), flush=True)
n = k ** m
x = np.random.default_rng(n).permutation(n) * ((1 - 1e-8) / n)
for rep in range(int(os.environ.get("REPS", "1"))):
    out, ms, level = s.sort(x)
    err = np.abs(out - np.sort(x)).max()
    print("k=%d M=%d N=%d ring 2^%d %s: %.3f s, level %d, max error %.3g (2^%.1f), guard bootstraps so far %d" %
          (k, m, n, logn, drivers, ms / 1e3, level, err, np.log2(err), s.bootstraps() if drivers == "ours" else -1), flush=True)
