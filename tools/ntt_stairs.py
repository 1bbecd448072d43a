#!/usr/bin/env python
"""ntt_fwd / ntt_inv isolated-kernel time against the number of limbs in the batch (ring 2^17): shows the wave structure of the
launch form selected by the environment (FHEB200_NTT_CLUSTER etc.).  python tools/ntt_stairs.py [tag]"""
import json
import sys

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402

ctx = F.Context(17, 39, 40, 60, batch=0, dnum=3, device=0)
out = {"tag": sys.argv[1] if len(sys.argv) > 1 else ""}
for name in ("ntt_fwd", "ntt_inv"):
    row = {}
    for ell in (1, 2, 4, 8, 9, 12, 16, 17, 18, 19, 20, 24, 32, 33, 36, 37, 41):
        ctx.bench_kernel(name, ell, 3)
        ms, _ = ctx.bench_kernel(name, ell, 30)
        row[ell] = round(1e3 * ms / 30, 1)
    out[name] = row
print(json.dumps(out), flush=True)
