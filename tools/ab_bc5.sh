#!/bin/bash
# tcgen05 base conversion against the mma.sync one: parity (ring 2^17 bit-exact tests + the ModDown pipeline) and kbench
OUT=gpurun_out/bc5
mkdir -p $OUT
for v in 0 1; do
  FHEB200_BC_TCGEN05=$v timeout 600 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_kernels_limb.py -m gpu -x -q -k "modup or rotate or moddown" > $OUT/pytest_$v.log 2>&1
  echo "FHEB200_BC_TCGEN05=$v parity exit $?" >> $OUT/kbench.txt
  FHEB200_BC_TCGEN05=$v timeout 300 python tools/kbench.py BC_TCGEN05=$v >> $OUT/kbench.txt 2>> $OUT/kbench.err
done
cat $OUT/kbench.txt; tail -n 12 $OUT/pytest_1.log
