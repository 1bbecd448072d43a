#!/bin/bash
# ncu launch list of ONE EvalBootstrap at the k-way parameter set (after the plain run has exited 0).
# Usage: gpurun --timeout 900 -- 'bash tools/boot_ncu.sh <tag>'
TAG=${1:-boot}
OUT=gpurun_out/$TAG
mkdir -p $OUT
REPS=1 FHEB200_PSTIME=1 FHEB200_BOOT_VERBOSE=1 timeout 200 python tools/boot_once.py 10 > $OUT/plain.log 2> $OUT/plain.err || { tail $OUT/plain.err; exit 1; }
cat $OUT/plain.log; grep -i "PS deg" $OUT/plain.err | tail -1
SKIP=$(grep -o 'timed bootstraps: [0-9]*' $OUT/plain.err | grep -o '[0-9]*$')
echo "ncu skip $SKIP"
REPS=1 timeout 700 ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c 1800 --csv \
    --log-file $OUT/boot_launches.csv python tools/boot_once.py 10 > $OUT/ncu.log 2>&1
python tools/ncu_summary.py $OUT/boot_launches.csv > $OUT/boot_launches_summary.txt; cat $OUT/boot_launches_summary.txt
