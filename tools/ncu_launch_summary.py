#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv): python tools/ncu_launch_summary.py file.csv"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
H = rows[0]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    if len(r) <= vi:
        continue
    name = re.sub(r"\(.*", "", r[ki]).replace("void fheb200::", "").replace("fheb200::", "")
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] in ("ns", "nsecond") else v * 1e3 if r[ui] in ("ms", "msecond") else v
    tot[name] += v
    cnt[name] += 1
total = sum(tot.values())
print("%-52s %8s %12s %9s %7s" % ("kernel", "n", "total_us", "avg_us", "share"))
for k in sorted(tot, key=lambda k: -tot[k]):
    print("%-52s %8d %12.1f %9.2f %7.3f" % (k[:52], cnt[k], tot[k], tot[k] / cnt[k], tot[k] / total))
print("%-52s %8d %12.1f" % ("total", sum(cnt.values()), total))
