#!/bin/bash
# like ab_build.sh but isolated-kernel timings only (for experiments whose results are wrong on purpose)
TAG=$1; SRC=$2; shift 2
OUT=gpurun_out/$TAG
mkdir -p $OUT
for f in "$@"; do
  touch fhe_sorting_b200/csrc/$SRC
  make -s -C fhe_sorting_b200 -j8 EXTRA="$f" > $OUT/build.log 2>&1 || { echo "build failed for $f"; tail -n 5 $OUT/build.log; continue; }
  timeout 300 python tools/kbench.py "$f" >> $OUT/kbench.txt 2>> $OUT/kbench.err
done
touch fhe_sorting_b200/csrc/$SRC
make -s -C fhe_sorting_b200 -j8 > $OUT/build.log 2>&1
cut -c1-120 $OUT/kbench.txt
