#!/usr/bin/env python
"""A few EvalRotate (and optionally EvalMult) calls at the bench parameter set inside a cudaProfilerStart/Stop bracket:
    ncu --profile-from-start off ... python tools/rot_once.py [nrot] [ell]
Prints the CUDA-event time per rotation when run without a profiler."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import fhe_sorting_b200 as F  # noqa: E402

nrot = int(sys.argv[1]) if len(sys.argv) > 1 else 2
ell = int(sys.argv[2]) if len(sys.argv) > 2 else 41
ctx = F.Context(17, 39, 40, 60, batch=0, dnum=3, device=0)
keys = ctx.keygen(7)
ctx.eval_rotate_keygen(keys, [1, 2, 3, 4])
ctx.eval_mult_keygen(keys)
vals = np.linspace(-1, 1, 1024)
ct = ctx.encrypt(keys, ctx.make_plaintext(vals, 1, 0, 1024))
while ct.limbs > ell:
    ct = ctx.mult(ct, 1.0)
    ct = ctx.rescale(ct) if ct.info()["nsd"] == 2 else ct
ctx.bench_rotations(ct, [1, 2, 3, 4], 3)
ms = ctx.bench_rotations(ct, [1, 2, 3, 4], 10)
print("rotate at %d limbs: %.1f us" % (ct.limbs, 1e3 * ms / 40), flush=True)
ctx.sync()
torch.cuda.profiler.start()
for i in range(nrot):
    r = ctx.rotate(ct, 1 + (i & 3))
m = ctx.mult(ct, ct)
ctx.sync()
torch.cuda.profiler.stop()
got = ctx.decrypt(keys, r, 1024)
print("err", float(np.abs(got - np.roll(vals, -(1 + ((nrot - 1) & 3)))).max()), flush=True)
