#!/usr/bin/env python
"""EvalBootstrap at the k-way parameter set (ring 2^17, depth 40, 59-bit scaling, 1024 slots, budget {5,5}):
latency and error per correction factor.  python tools/boot_once.py [corrections...]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fhe_sorting_b200 as F
logn = int(os.environ.get("LOGN", "17"))
slots = int(os.environ.get("SLOTS", "1024"))
budget = (int(os.environ.get("BUDGET", "5")),) * 2
for c in [int(a) for a in sys.argv[1:]] or [12]:
    os.environ["FHEB200_BOOT_CORRECTION"] = str(c)
    ctx = F.Context(logn, 40, 59, 60, batch=slots)
    keys = ctx.keygen(5)
    ctx.eval_mult_keygen(keys)
    ctx.bootstrap_setup(budget, slots)
    ctx.bootstrap_keygen(keys, slots)
    z = np.random.default_rng(1).uniform(0, 1, slots)
    ct = ctx.encrypt(keys, ctx.make_plaintext(z, 1, 0, slots))
    while ct.level < ctx.depth - 1:
        ct = ctx.mult(ct, 1.0)
    out = ctx.bootstrap(ct)  # first call encodes the diagonals
    ctx.sync()
    print("launches before the timed bootstraps: %d" % F.launch_count(), file=sys.stderr, flush=True)
    t0 = time.time()
    reps = int(os.environ.get("REPS", "5"))
    for _ in range(reps):
        out = ctx.bootstrap(ct)
    ctx.sync()
    ms = (time.time() - t0) / reps * 1e3
    err = np.abs(ctx.decrypt(keys, out, slots) - z)
    print("correction %d: %.1f ms per EvalBootstrap, output level %d, max error 2^%.1f, mean 2^%.1f, key bytes %.1f GiB" %
          (c, ms, out.level, np.log2(err.max()), np.log2(err.mean()), ctx.key_bytes() / 2**30), flush=True)
    if os.environ.get("FHEB200_BOOT_VERBOSE"):
        break
    del ctx, keys, ct, out
