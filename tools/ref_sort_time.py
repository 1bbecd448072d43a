#!/usr/bin/env python
"""DirectSort N through the reference's own drivers (OpenMP host threads, one CUDA stream each) against this repository's
drivers (one host thread, forked streams): python tools/ref_sort_time.py N"""
import ctypes as C
import os
import sys
import numpy as np
sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
n = int(sys.argv[1])
x = uniform_gap_input(n)
for name, lib in (("own", None), ("reference", "oracle/_ref/libref_sort.so")):
    L = None
    if lib:
        L = C.CDLL(os.path.abspath(lib)); L.fhesort_last_error.restype = C.c_char_p
    s = DirectSortSession(n, 17, seed=1000 + n, lib=L)
    for rep in range(3):
        out, ms, level = s.sort(x)
    print("%s drivers N=%d: %.1f ms err %.2e level %d (OMP_NUM_THREADS=%s)" % (name, n, ms, np.abs(out - np.sort(x)).max(), level, os.environ.get("OMP_NUM_THREADS")), flush=True)
    s.close()
