#!/bin/bash
# A/B of kernel variants selected by one environment variable, at the bench parameter set, with the ring-2^17 parity tests
# run under each value.  usage: gpurun -- 'bash tools/ab_env.sh <tag> <VAR> <v1> <v2> ...'
TAG=$1; VAR=$2; shift 2
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in "$@"; do
  env $VAR=$v timeout 300 python tools/kbench.py $VAR=$v >> $OUT/kbench.txt 2>> $OUT/kbench.err
  env $VAR=$v timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "ntt or modup or rotate" > $OUT/pytest_$v.log 2>&1
  echo "$VAR=$v parity exit $?" >> $OUT/kbench.txt
done
cat $OUT/kbench.txt
