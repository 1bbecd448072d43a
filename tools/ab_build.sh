#!/bin/bash
# A/B of a compile-time switch of one source file, built on the GPU box: bash tools/ab_build.sh <tag> <file.cu> <flag1> <flag2> ...
# (e.g. ntt.cu -DNTT_CTAS_X=4 -DNTT_CTAS_X=5); the default build is restored at the end.
TAG=$1; SRC=$2; shift 2
OUT=gpurun_out/$TAG
mkdir -p $OUT
for f in "$@"; do
  touch fhe_sorting_b200/csrc/$SRC
  make -s -C fhe_sorting_b200 -j8 EXTRA="$f" > $OUT/build.log 2>&1 || { echo "build failed for $f"; tail -n 5 $OUT/build.log; continue; }
  echo "== $f" >> $OUT/kbench.txt
  timeout 300 python tools/kbench.py "$f" >> $OUT/kbench.txt 2>> $OUT/kbench.err
  timeout 300 python tools/sort_once.py 256 3 2>&1 | tail -n 1 >> $OUT/kbench.txt
  timeout 300 python tools/sort_once.py 1024 2 2>&1 | tail -n 1 >> $OUT/kbench.txt
done
touch fhe_sorting_b200/csrc/$SRC
make -s -C fhe_sorting_b200 -j8 > $OUT/build.log 2>&1
cat $OUT/kbench.txt
