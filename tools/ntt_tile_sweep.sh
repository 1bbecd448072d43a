#!/bin/bash
# rebuild libfheb200.so with different NTT tile sizes on the GPU box and time the kernels
for t in 12 11 10; do
  rm -f fhe_sorting_b200/build/ntt.cu.o
  make -s -C fhe_sorting_b200 EXTRA="-DNTT_TILE_LOG_V=$t" -j16 > gpurun_out/sweep_build_$t.log 2>&1
  python tools/kbench.py tile$t 2>&1 | tail -1 | tee -a gpurun_out/ntt_tile_sweep.txt
done
