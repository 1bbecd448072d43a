#!/bin/bash
# ncu evidence for the bench command (one gpurun call): launch list of the timed region + --set full of one rotation's kernels,
# both only after the plain command exited 0; summaries made on the box.
OUT=gpurun_out/r02ncu
mkdir -p $OUT
timeout 300 python bench.py --skip-sort --skip-micro --steps 2 --warmup 3 > $OUT/plain.log 2> $OUT/plain.err || exit 1
SKIP=$(grep -o 'timed region: [0-9]*' $OUT/plain.err | grep -o '[0-9]*$')
echo "ncu skip $SKIP"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c 440 --csv \
    --log-file $OUT/launches.csv python bench.py --skip-sort --skip-micro --steps 2 --warmup 3 > $OUT/ncu_launches.log 2>&1
python tools/ncu_launch_summary.py $OUT/launches.csv > $OUT/launches_summary.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -s $SKIP -c 14 \
    -o $OUT/prof python bench.py --skip-sort --skip-micro --steps 2 --warmup 3 > $OUT/ncu_full.log 2>&1
python tools/ncu_raw.py $OUT/prof.ncu-rep > $OUT/ncu_full_bench.txt 2>&1
rm -f $OUT/prof.ncu-rep
cat $OUT/launches_summary.txt; cut -c1-400 $OUT/ncu_full_bench.txt
