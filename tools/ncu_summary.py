#!/usr/bin/env python
"""Summarise an ncu launch list (gpu__time_duration.sum CSV) per kernel: count, total, average, share."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
ki, vi, ui = H.index('Kernel Name'), H.index('Metric Value'), H.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(',', ''))
    v = v / 1e3 if r[ui] == 'ns' else v * 1e3 if r[ui] == 'ms' else v
    name = r[ki].split('(')[0]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print("%-40s %6s %12s %10s %7s" % ("kernel", "n", "total_us", "avg_us", "share"))
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-40s %6d %12.1f %10.1f %7.3f" % (k, c, t, t / c, t / tot))
print("%-40s %6d %12.1f" % ("total", sum(v[0] for v in agg.values()), tot))
