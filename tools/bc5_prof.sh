#!/bin/bash
# ncu --set full of the tcgen05 base conversion (one gpurun call); summaries made on the box
OUT=gpurun_out/bc5prof
mkdir -p $OUT
export FHEB200_BC_TCGEN05=1
cat > /tmp/bc_one.py <<'PY'
import sys
sys.path.insert(0, ".")
import fhe_sorting_b200 as F
ctx = F.Context(17, 39, 40, 60, batch=0, dnum=3, device=0)
for name in ("modup_conv", "moddown_conv"):
    ms, _ = ctx.bench_kernel(name, 41, 3)
    print(name, round(1e3 * ms / 3, 1), "us")
PY
python /tmp/bc_one.py > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'bc_convert_t5' -c 2 -o $OUT/bc python /tmp/bc_one.py > $OUT/ncu.log 2>&1
python tools/ncu_raw.py $OUT/bc.ncu-rep > $OUT/ncu_full_bc5.txt 2>&1
ncu -i $OUT/bc.ncu-rep --page source --csv > $OUT/source_raw.csv 2> $OUT/source_err.txt
head -c 3000 $OUT/source_raw.csv > $OUT/source_head.txt
python tools/ncu_source_top.py 60 < $OUT/source_raw.csv > $OUT/source_top.txt 2>&1
gzip -f $OUT/source_raw.csv
rm -f $OUT/bc.ncu-rep
cat $OUT/plain.log; tail -n 3 $OUT/ncu.log; cut -c1-900 $OUT/ncu_full_bc5.txt | head -n 4; head -n 70 $OUT/source_top.txt
