#!/bin/bash
OUT=gpurun_out/merged
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_scheme.py tests/test_gpu_fullsize.py tests/test_gpu_sort.py tests/test_gpu_bootstrap.py tests/test_gpu_mehp24.py tests/test_gpu_cpu_crosscheck.py -m gpu -x -q -k "not kway" > $OUT/pytest.log 2>&1
echo "parity exit $?" > $OUT/summary.txt
for v in 0 1; do
  echo "== FHEB200_MERGED_RESCALE=$v" >> $OUT/summary.txt
  FHEB200_MERGED_RESCALE=$v timeout 300 python tools/sort_once.py 256 3 2>&1 | tail -n 2 >> $OUT/summary.txt
  FHEB200_MERGED_RESCALE=$v timeout 300 python tools/sort_once.py 1024 3 2>&1 | tail -n 2 >> $OUT/summary.txt
done
cat $OUT/summary.txt; tail -n 25 $OUT/pytest.log
