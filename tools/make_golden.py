#!/usr/bin/env python
"""Golden fixtures from the reference itself: oracle/_ref/gen/*.h are written by the reference's own, unmodified
utils/generate_cheb_coeffs.cpp and utils/generate_cheb_doubled_coeffs.cpp (compiled by `make -C oracle ref`, which needs
/root/reference).  This script copies the small tables into tests/golden/ so that the comparison also runs where the
reference is absent.  python tools/make_golden.py"""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "oracle", "_ref", "gen")
out = {}
for fname, var, key in (("generated_coeffs.h", "generatedCoefficients", "sinc"),
                        ("generated_doubled_sinc_coeffs.h", "generatedDoubledSincCoefficients", "doubled_sinc")):
    txt = open(os.path.join(GEN, fname)).read()
    out[key] = {}
    for n in (4, 8, 16, 32, 64, 128):
        m = re.search(r"%s_%d = \{(.*?)\};" % (var, n), txt, re.S)
        out[key][str(n)] = [float(v) for v in m.group(1).replace("\n", " ").split(",") if v.strip()]
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "chebyshev_tables.json"), "w"))
print({k: {n: len(v) for n, v in t.items()} for k, t in out.items()})

# k-way network geometry from the reference's own src/k-way/Masking.cpp (compiled into oracle/_ref/libref_sort.so)
import sys
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from fhe_sorting_b200 import sort as S  # noqa: E402
import refdrivers  # noqa: E402

REF = refdrivers.library()

kw = {}
for k, m in ((2, 2), (2, 3), (2, 5), (3, 2), (3, 3), (5, 2), (5, 3)):
    rows = []
    for st in range(S.kway_num_stages(k, m)):
        geo, idx = S.kway_stage(k, m, st, k ** m, lib=REF)
        rows.append({"geometry": list(geo), "indices": idx.tolist()})
    kw["%d^%d" % (k, m)] = rows
json.dump(kw, open(os.path.join(ROOT, "tests", "golden", "kway_stages.json"), "w"), separators=(",", ":"))
print({k: len(v) for k, v in kw.items()})
