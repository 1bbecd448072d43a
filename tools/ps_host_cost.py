#!/usr/bin/env python
"""Host-side cost of one degree-6510 Paterson-Stockmeyer evaluation: the same call sequence on a 2^12 ring, where every
kernel is a few microseconds, so the elapsed time is what the launching thread (or per-launch latency) costs."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402
from fhe_sorting_b200.sort import doubled_sinc_coefficients  # noqa: E402

coeffs = doubled_sinc_coefficients(1024)
for logn in (12, 17):
    ctx = F.Context(logn, 39, 40, 60, batch=8)
    keys = ctx.keygen(5)
    ctx.eval_mult_keygen(keys)
    ct = ctx.encrypt(keys, ctx.make_plaintext(np.linspace(-0.9, 0.4, 8), 1, 0, 8))
    for _ in range(22):  # burn levels as the sort has by then
        ct = ctx.mult(ct, 1.0)
        ct = ctx.rescale(ct) if ct.info()["nsd"] == 2 else ct
    ctx.sync()
    for rep in range(3):
        l0 = F.launch_count()
        t0 = time.perf_counter()
        r = ctx.chebyshev(ct, coeffs, -1, 1)
        t1 = time.perf_counter()
        ctx.sync()
        t2 = time.perf_counter()
        print("ring 2^%d ell %d: PS enqueued in %.1f ms, done after %.1f ms, %d launches" %
              (logn, ct.limbs, 1e3 * (t1 - t0), 1e3 * (t2 - t0), F.launch_count() - l0), flush=True)
