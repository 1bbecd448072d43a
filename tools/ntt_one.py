#!/usr/bin/env python
"""A few forward + inverse NTT launches of `ell` limbs at ring 2^17 (for ncu): python tools/ntt_one.py [ell]"""
import sys

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402

ell = int(sys.argv[1]) if len(sys.argv) > 1 else 16
ctx = F.Context(17, 39, 40, 60, batch=0, dnum=3, device=0)
for name in ("ntt_fwd", "ntt_inv"):
    ms, _ = ctx.bench_kernel(name, ell, 4)
    print(name, ell, "limbs", round(1e3 * ms / 4, 1), "us")
