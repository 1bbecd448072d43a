#!/bin/bash
# kernel-time breakdown of the first COUNT launches of rotationIndexCheckN (one PS evaluation is ~5700 launches at N=1024)
N=${1:-1024}; COUNT=${2:-6000}
python tools/place_once.py $N > gpurun_out/place_$N.txt 2> gpurun_out/place_$N.err || exit 1
cat gpurun_out/place_$N.txt
SKIP=$(grep -o 'at launch: [0-9]*' gpurun_out/place_$N.err | grep -o '[0-9]*$')
echo "skip $SKIP"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c $COUNT --csv --log-file /tmp/place_launches.csv \
    python tools/place_once.py $N > gpurun_out/place_ncu.log 2>&1
python tools/ncu_summary.py /tmp/place_launches.csv > gpurun_out/place_kernels_$N.txt
head -45 gpurun_out/place_kernels_$N.txt; tail -2 gpurun_out/place_kernels_$N.txt
