#!/usr/bin/env python
"""Encrypted DirectSort latency at the reference's parameters (ring 2^17): python tools/sort_latency.py 16 64 256 ..."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input  # noqa: E402
import fhe_sorting_b200 as F  # noqa: E402

for n in [int(a) for a in sys.argv[1:]] or [16]:
    t0 = time.time()
    s = DirectSortSession(n, 17, seed=1000 + n)
    setup = time.time() - t0
    x = uniform_gap_input(n)
    res = []
    wall = []
    for rep in range(4):
        l0 = F.launch_count()
        t1 = time.perf_counter()
        out, ms, level = s.sort(x)
        wall.append(round(1e3 * (time.perf_counter() - t1), 1))
        res.append(ms)
        launches = F.launch_count() - l0
    err = float(np.abs(out - np.sort(x)).max())
    print(json.dumps({"N": n, "sort_ms": res, "wall_ms": wall, "max_err": err, "level": level, "depth": s.mult_depth,
                      "rot_keys": s.num_rotation_keys, "setup_s": round(setup, 2), "kernel_launches": launches}), flush=True)
    s.close()
