import sys, numpy as np
sys.path.insert(0, ".")
from fhe_sorting_b200.sort import DirectSortSession, uniform_gap_input
n = int(sys.argv[1]); streams = int(sys.argv[2])
s = DirectSortSession(n, 17, seed=1000 + n)
if streams: s.set_streams(streams)
x = uniform_gap_input(n)
for rep in range(2):
    out, ms, level = s.sort(x)
    print("N=%d streams=%d rep %d: %.1f ms err %.2e" % (n, streams, rep, ms, np.abs(out - np.sort(x)).max()), flush=True)
