#!/bin/bash
# GPU-busy time of one encrypted sort: ncu launch list (durations only), summarised per kernel on the box
N=${1:-256}
python tools/sort_once.py $N 3 > gpurun_out/sort_once_$N.txt 2>&1 || exit 1
cat gpurun_out/sort_once_$N.txt
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file /tmp/sort_launches_$N.csv \
    python tools/sort_once.py $N 1 > gpurun_out/sort_ncu_$N.log 2>&1
python tools/ncu_summary.py /tmp/sort_launches_$N.csv > gpurun_out/sort_kernels_$N.txt
head -40 gpurun_out/sort_kernels_$N.txt
