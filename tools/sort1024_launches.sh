#!/bin/bash
# launch list of one DirectSort N = 1024 (ncu, gpu__time_duration only), summarised on the box
OUT=gpurun_out/sort1024
mkdir -p $OUT
python tools/sort_once.py 1024 2 > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $OUT/launches.csv \
    python tools/sort_once.py 1024 2 > $OUT/ncu.log 2>&1
python tools/ncu_launch_summary.py $OUT/launches.csv > $OUT/sort1024_kernels.txt 2>&1
gzip -f $OUT/launches.csv
tail -n 2 $OUT/plain.log; head -n 40 $OUT/sort1024_kernels.txt
