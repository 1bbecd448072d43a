#!/bin/bash
# tcgen05 leaf sums against lc_tc_ring: limb-level parity (all variants vs exact integers) and timing
OUT=gpurun_out/lc5
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_kernels_limb.py -m gpu -x -q -k "leaf_sums" > $OUT/pytest.log 2>&1
echo "parity exit $?" > $OUT/summary.txt
for v in 0 1; do
  FHEB200_LC_TCGEN05=$v timeout 300 python tools/lincomb_bench.py 2>&1 | grep -v checksum | sed "s/^/LC_TCGEN05=$v /" >> $OUT/summary.txt
done
cat $OUT/summary.txt; tail -n 15 $OUT/pytest.log
