#!/bin/bash
# A/B of one environment variable on the sorts (+ the ring-2^17 parity tests under each value).
# usage: gpurun -- 'bash tools/ab_sort.sh <tag> <VAR> <v1> <v2> ...'
TAG=$1; VAR=$2; shift 2
OUT=gpurun_out/$TAG
mkdir -p $OUT
for v in "$@"; do
  echo "== $VAR=$v" >> $OUT/sorts.txt
  env $VAR=$v timeout 600 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_scheme.py tests/test_gpu_primitives.py -m gpu -x -q -k "not kway" > $OUT/pytest_$v.log 2>&1
  echo "parity exit $?" >> $OUT/sorts.txt
  env $VAR=$v timeout 300 python tools/sort_once.py 256 3 2>&1 | tail -n 2 >> $OUT/sorts.txt
  env $VAR=$v timeout 300 python tools/sort_once.py 1024 3 2>&1 | tail -n 2 >> $OUT/sorts.txt
done
cat $OUT/sorts.txt
