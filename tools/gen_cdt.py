#!/usr/bin/env python3
"""Generate the 63-bit cumulative table of the folded discrete Gaussian (sigma = 3.19)
shared verbatim by oracle/ckks_oracle.cpp and fhe_sorting_b200/csrc (sampler spec in DESIGN.md).
CDT[k] = floor(2^63 * P(|X| <= k)); a 63-bit uniform r maps to |x| = #{k : r >= CDT[k]}."""
from decimal import Decimal, getcontext
getcontext().prec = 80
sigma = Decimal("3.19")
two_s2 = 2 * sigma * sigma
def rho(k):
    return (-(Decimal(k) * Decimal(k)) / two_s2).exp()
K = 64
w = [rho(k) for k in range(K)]
S = w[0] + 2 * sum(w[1:])
cum = Decimal(0)
out = []
for k in range(K):
    cum += (w[k] if k == 0 else 2 * w[k]) / S
    v = int(cum * (1 << 63))
    if v >= (1 << 63) - 1:
        break
    out.append(v)
print(len(out))
for i in range(0, len(out), 4):
    print("    " + ", ".join("0x%016xULL" % v for v in out[i:i + 4]) + ",")
