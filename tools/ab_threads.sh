#!/bin/bash
OUT=gpurun_out/threads
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_sort.py tests/test_gpu_fullsize.py tests/test_gpu_mehp24.py tests/test_gpu_cpu_crosscheck.py -m gpu -x -q -k "not kway" > $OUT/pytest.log 2>&1
echo "parity exit $?" > $OUT/summary.txt
for v in 0 1; do
  echo "== FHESORT_THREADS=$v" >> $OUT/summary.txt
  FHESORT_THREADS=$v timeout 300 python tools/sort_once.py 1024 4 2>&1 | grep "N=" | tail -n 3 >> $OUT/summary.txt
  FHESORT_THREADS=$v timeout 300 python tools/sort_once.py 512 3 2>&1 | grep "N=" | tail -n 2 >> $OUT/summary.txt
done
for s in 2 8; do echo "== threads, streams $s" >> $OUT/summary.txt; FHESORT_STREAMS=$s timeout 300 python tools/sort_once.py 1024 3 2>&1 | grep "N=" | tail -n 2 >> $OUT/summary.txt; done
cat $OUT/summary.txt; tail -n 15 $OUT/pytest.log
