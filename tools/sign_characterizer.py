#!/usr/bin/env python
"""Precision / depth / time sweep of the sign functions on the B200 engine: the procedure of the reference's
utils/sign_characterizer.cpp:60-105 (context of :26-38: depth 40, scaling 50 bits, batch 8), over the reference-shaped
C++ sign() dispatcher (host/sign.cpp) through libfhesort.so.

For each function: starting from step = 0.1, evaluate sign(+-step) on a fresh encryption; while the error against the
exact sign stays below the threshold, divide step by 10.  Reported: the last level, the working precision (10 x the first
failing step; 0 = never failed within the value budget) and the time of the last evaluation.

    python tools/sign_characterizer.py [--threshold 0.01] [--max-values 100] [--header out.h]
"""
import argparse
import sys

import numpy as np

sys.path.insert(0, ".")
from fhe_sorting_b200 import sort as S  # noqa: E402

FUNCS = [("NaiveDiscrete", "NaiveDiscrete", (0, 0, 0)), ("Tanh", "Tanh", (0, 0, 0)), ("SignumPolycircuit", "SignumPolycircuit", (0, 0, 0)),
         ("CompositeSign_2_2", "CompositeSign", (4, 2, 2)), ("CompositeSign_3_3", "CompositeSign", (4, 3, 3))]


def characterize(sess, func, cfg, threshold, max_values):
    tested, step, depth, ms, working = 3, 0.1, 0, 0.0, 0.0       # the reference seeds its list with -1, 0, 1
    while tested < max_values:
        stop = False
        for value in (-step, step):
            if tested >= max_values:
                break
            out, depth, ms = sess.sign(func, np.array([value]), *cfg)
            expected = 0.0 if func == "NaiveDiscrete" and value < 0 else float(np.sign(value))
            if abs(out[0] - expected) > threshold:
                stop, working = True, step * 10
                break
            tested += 1
        if stop:
            break
        step /= 10
    return depth, working, ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threshold", type=float, default=0.01)
    ap.add_argument("--max-values", type=int, default=100)
    ap.add_argument("--header", default="")
    ap.add_argument("--log-ring-dim", type=int, default=16)      # OpenFHE picks 2^16 for depth 40 / scaling 50 at 128 bits
    a = ap.parse_args()
    sess = S.ToolkitSession(a.log_ring_dim, 40, 50, 8, seed=1)
    rows = []
    for name, func, cfg in FUNCS:
        depth, working, ms = characterize(sess, func, cfg, a.threshold, a.max_values)
        rows.append((name, depth, working, ms))
        print("%-20s depth %2d   working precision %-8g   %.2f ms" % (name, depth, working, ms), flush=True)
    if a.header:
        with open(a.header, "w") as f:
            f.write("// Precision Threshold: %g\n// Max Test Values: %d\n#ifdef GET_SIGN_FUNCTION_METRIC\n" % (a.threshold, a.max_values))
            for name, depth, working, ms in rows:
                f.write("GET_SIGN_FUNCTION_METRIC(%s, %d, %.10g, %g)\n" % (name, depth, working, ms))
            f.write("#endif // GET_SIGN_FUNCTION_METRIC\n")
    sess.close()


if __name__ == "__main__":
    main()
