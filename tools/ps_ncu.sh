#!/bin/bash
# kernel-time breakdown of one in-sort Paterson-Stockmeyer evaluation (N=1024, the levels it really runs at)
FHEB200_PSTIME=1 FHESORT_STREAMS=1 python tools/sort_once.py 1024 1 2> gpurun_out/ps_lines.txt > /dev/null || exit 1
A=$(grep -o 'launch [0-9]*' gpurun_out/ps_lines.txt | sed -n 2p | grep -o '[0-9]*')
B=$(grep -o 'launch [0-9]*' gpurun_out/ps_lines.txt | sed -n 3p | grep -o '[0-9]*')
echo "third PS call: launches $A .. $B"
FHESORT_STREAMS=1 timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s $A -c $((B-A)) --csv --log-file /tmp/ps_launches.csv \
    python tools/sort_once.py 1024 1 > gpurun_out/ps_ncu.log 2>&1
python tools/ncu_summary.py /tmp/ps_launches.csv > gpurun_out/ps_kernels_1024.txt
cat gpurun_out/ps_kernels_1024.txt | head -40
