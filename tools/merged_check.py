#!/usr/bin/env python
"""relinearise+rescale as one division against the two-step form at ring 2^17: coefficient-domain difference per limb"""
import sys
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import fhe_sorting_b200 as F
ctx = F.Context(17, 39, 40, 60)
keys = ctx.keygen(1170)
ctx.eval_mult_keygen(keys)
rng = np.random.default_rng(8)
va, vb = rng.uniform(-1, 1, 1024), rng.uniform(-1, 1, 1024)
a = ctx.encrypt(keys, ctx.make_plaintext(va, 1, 0, 1024))
b = ctx.encrypt(keys, ctx.make_plaintext(vb, 1, 0, 1024))
two, one = ctx.rescale(ctx.mult(a, b)), ctx.mult_rescaled(a, b)
print(one.info(), two.info())
print("err one", np.abs(ctx.decrypt(keys, one, 1024) - va * vb).max(), "err two", np.abs(ctx.decrypt(keys, two, 1024) - va * vb).max())
x0, x1 = ctx.ct_export(one)
y0, y1 = ctx.ct_export(two)
for i in (0, 1, 2, 20, 38):
    q = ctx.moduli[i]
    for nm, x, y in (("c0", x0, y0), ("c1", x1, y1)):
        d = (ctx.raw_ntt(i, x[i], inverse=True).astype(object) - ctx.raw_ntt(i, y[i], inverse=True).astype(object)) % q
        d = np.array([int(v) - q if int(v) > q // 2 else int(v) for v in d], dtype=np.float64)
        print("limb", i, nm, "max |diff|", np.abs(d).max(), "first", d[:6])
