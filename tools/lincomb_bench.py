#!/usr/bin/env python
"""Leaf-sum kernel (k_lincomb, 63 inputs -> 4 outputs over 2*ell limbs) timing per variant: FHEB200_LINCOMB=1|2"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fhe_sorting_b200 as F
ctx = F.Context(17, 39, 40, 60)
for ell in (8, 14, 20):
    ms, by = ctx.bench_kernel("lincomb", ell, 10)
    print("FHEB200_LINCOMB=%s ell %d: %.1f us per launch, %.0f GB/s algorithmic" % (os.environ.get("FHEB200_LINCOMB", "default"), ell, ms / 10 * 1e3, by / (ms / 10 * 1e-3) / 1e9), flush=True)
