#!/usr/bin/env python
"""Top SASS instructions of a kernel by warp-stall samples, from `ncu -i rep --page source --csv --kernel-name ...` on stdin.
python tools/ncu_source_top.py [count]"""
import csv
import sys

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rows = list(csv.reader(sys.stdin))
hdr = next(i for i, r in enumerate(rows) if r and "Source" in r and "# Samples" in r)
H = rows[hdr]
si, ii, smp = H.index("Source"), H.index("Instructions Executed"), H.index("# Samples")
body = [r for r in rows[hdr + 1:] if len(r) > smp and (r[smp] or "0").isdigit()]
total = sum(int(r[smp] or 0) for r in body)
print("instructions %d, samples %d" % (len(body), total))
order = sorted(range(len(body)), key=lambda i: -int(body[i][smp] or 0))[:n]
for i in sorted(order):
    r = body[i]
    print("%5d %6.2f%%  exec %8s  %s" % (i, 100.0 * int(r[smp] or 0) / max(total, 1), r[ii], r[si].strip()))
