#!/bin/bash
OUT=gpurun_out/lc5prof
mkdir -p $OUT
export FHEB200_LC_TCGEN05=1
cat > /tmp/lc_one.py <<'PY'
import sys
sys.path.insert(0, ".")
import fhe_sorting_b200 as F
ctx = F.Context(17, 39, 40, 60)
ms, by = ctx.bench_kernel("lincomb", 14, 3)
print("lincomb", round(1e3 * ms / 3, 1), "us")
PY
python /tmp/lc_one.py > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'lc_tc_t5' -c 2 -o $OUT/lc python /tmp/lc_one.py > $OUT/ncu.log 2>&1
python tools/ncu_raw.py $OUT/lc.ncu-rep > $OUT/ncu_full_lc5.txt 2>&1
ncu -i $OUT/lc.ncu-rep --page source --csv 2>/dev/null > $OUT/source_raw.csv; python tools/ncu_source_top.py 40 < $OUT/source_raw.csv > $OUT/source_top.txt 2>&1; gzip -f $OUT/source_raw.csv
ncu -i $OUT/lc.ncu-rep --page details 2>/dev/null | grep -E "Bank|bank|Wavefront|Excessive|L1/TEX Hit|Shared|Throughput|Stall|stall" | head -n 60 > $OUT/details.txt
rm -f $OUT/lc.ncu-rep
cat $OUT/plain.log; tail -n 2 $OUT/ncu.log; cut -c1-900 $OUT/ncu_full_lc5.txt | head -n 3; head -n 5 $OUT/source_top.txt
