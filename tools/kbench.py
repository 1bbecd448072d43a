#!/usr/bin/env python
"""Isolated kernel timings at the bench parameter set (ring 2^17, 41 limbs): python tools/kbench.py [tag]"""
import json
import sys

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402

ctx = F.Context(17, 39, 40, 60, batch=0, dnum=3, device=0)
keys = ctx.keygen(7)
ctx.eval_rotate_keygen(keys, [1, 2, 3, 4])
ctx.eval_mult_keygen(keys)
import numpy as np  # noqa: E402
ct = ctx.encrypt(keys, ctx.make_plaintext(np.linspace(-1, 1, 1024), 1, 0, 1024))
out = {"tag": sys.argv[1] if len(sys.argv) > 1 else ""}
for name in ("ntt_fwd", "ntt_inv", "inner_product", "modup_conv", "moddown_conv", "modup_conv_scalar", "moddown_conv_scalar"):
    ctx.bench_kernel(name, 41, 5)
    ms, b = ctx.bench_kernel(name, 41, 50)
    out[name + "_us"] = round(1e3 * ms / 50, 2)
ctx.bench_rotations(ct, [1, 2, 3, 4], 5)
ms = ctx.bench_rotations(ct, [1, 2, 3, 4], 50)
out["rotate_us"] = round(1e3 * ms / 200, 1)
r = ctx.rotate(ct, 1)
got = ctx.decrypt(keys, r, 1024)
out["rotate_err"] = float(np.abs(got - np.roll(np.linspace(-1, 1, 1024), -1)).max())
print(json.dumps(out), flush=True)
