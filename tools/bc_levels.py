#!/usr/bin/env python
"""ModUp / ModDown base conversion (isolated kernels) across levels: python tools/bc_levels.py [tag]"""
import json
import sys

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402

ctx = F.Context(17, 39, 40, 60, batch=0, dnum=3, device=0)
out = {"tag": sys.argv[1] if len(sys.argv) > 1 else ""}
for name in ("modup_conv", "moddown_conv"):
    row = {}
    for ell in (41, 36, 30, 28, 24, 20, 16, 14, 10, 6, 2):
        ctx.bench_kernel(name, ell, 3)
        ms, _ = ctx.bench_kernel(name, ell, 30)
        row[ell] = round(1e3 * ms / 30, 1)
    out[name] = row
print(json.dumps(out), flush=True)
