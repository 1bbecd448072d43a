#!/usr/bin/env python
"""Kernels outside the rotation for an ncu capture (cudaProfilerStart/Stop bracket): one degree-6510 Paterson-Stockmeyer
evaluation at 17 limbs (lc_tc_ring leaf sums, ew_tensor, ew_rescale_*), a 65536-slot encode (enc_*), plaintext product, and one
EvalBootstrap at the k-way parameters (ks_dh_accumulate).
    ncu --profile-from-start off ... python tools/prof_misc.py"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402

ctx = F.Context(17, 39, 40, 60, batch=65536, dnum=3, device=0)
keys = ctx.keygen(7)
ctx.eval_mult_keygen(keys)
rng = np.random.default_rng(1)
vals = rng.uniform(-1, 1, 65536)
ct = ctx.encrypt(keys, ctx.make_plaintext(vals, 1, 0, 65536))
while ct.limbs > 17:
    if ct.info()["nsd"] == 1:
        ct = ctx.mult(ct, 1.0)
    ct = ctx.rescale(ct)
coeffs = rng.uniform(-1, 1, 6511) / np.arange(1, 6512)
ctx.chebyshev(ct, coeffs, -1.0, 1.0)
bctx = F.Context(17, 40, 59, 60, batch=1024, device=0)
bkeys = bctx.keygen(7)
bctx.eval_mult_keygen(bkeys)
bctx.bootstrap_setup((5, 5), 1024)
bctx.bootstrap_keygen(bkeys, 1024)
bc = bctx.encrypt(bkeys, bctx.make_plaintext(vals[:1024], 1, 0, 1024))
while bc.level < bctx.depth - 1:
    bc = bctx.mult(bc, 1.0)
bctx.bootstrap(bc)
ctx.sync(), bctx.sync()
torch.cuda.profiler.start()
y = ctx.chebyshev(ct, coeffs, -1.0, 1.0)
pt = ctx.make_plaintext(rng.uniform(-1, 1, 65536), 1, 0, 65536)
z = ctx.mult(ct, pt)
b = bctx.bootstrap(bc)
ctx.sync(), bctx.sync()
torch.cuda.profiler.stop()
print("ok", y.level, z.level, b.level, flush=True)
