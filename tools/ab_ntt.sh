#!/bin/bash
# A/B of the NTT launch forms at the bench parameter set + the GPU test-suite.  usage: gpurun -- 'bash tools/ab_ntt.sh <tag>'
TAG=${1:-ab}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for f in 0 1; do
  FHEB200_NTT_FUSED=$f timeout 300 python tools/kbench.py fused$f >> $OUT/kbench.txt 2>> $OUT/kbench.err
done
cat $OUT/kbench.txt
timeout 1800 python -m pytest tests -m gpu -x -q --durations=15 > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -30 $OUT/pytest_gpu.log
