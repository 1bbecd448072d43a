#!/bin/bash
# ncu --set full of the cluster NTT kernels (one gpurun call); summaries are made on the box, the report does not travel
OUT=gpurun_out/clprof
mkdir -p $OUT
export FHEB200_NTT_CLUSTER=1
python tools/ntt_one.py 16 > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'ntt_(fwd|inv)_cluster' -c 4 -o $OUT/cl python tools/ntt_one.py 16 > $OUT/ncu.log 2>&1
python tools/ncu_raw.py $OUT/cl.ncu-rep > $OUT/ncu_full_cluster.txt 2>&1
ncu -i $OUT/cl.ncu-rep --page source --csv --kernel-name regex:ntt_fwd_cluster 2>/dev/null | python tools/ncu_source_top.py 40 > $OUT/source_top_fwd.txt
ncu -i $OUT/cl.ncu-rep --page source --csv --kernel-name regex:ntt_inv_cluster 2>/dev/null | python tools/ncu_source_top.py 40 > $OUT/source_top_inv.txt
ncu -i $OUT/cl.ncu-rep --page details --kernel-name regex:ntt_fwd_cluster 2>/dev/null | head -n 400 > $OUT/details_fwd.txt
rm -f $OUT/cl.ncu-rep
cat $OUT/plain.log; tail -n 3 $OUT/ncu.log; cat $OUT/ncu_full_cluster.txt | head -n 12
