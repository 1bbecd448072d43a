#!/bin/bash
# rebuild baseconv.cu with different tilings on the GPU box and time the conversion kernels
for cfg in "2 4" "1 4" "1 8" "2 8"; do
  set -- $cfg
  rm -f fhe_sorting_b200/build/baseconv.cu.o
  make -s -C fhe_sorting_b200 EXTRA="-DBC_MT_V=$1 -DBC_WARPS_V=$2" -j16 > gpurun_out/bc_build.log 2>&1
  grep -A2 "bc_convertILi4" fhe_sorting_b200/build/baseconv.cu.ptxas.log | grep -o "Used [0-9]* registers"
  python tools/kbench.py "mt$1_w$2" 2>&1 | tail -1 | tee -a gpurun_out/bc_sweep.txt
done
