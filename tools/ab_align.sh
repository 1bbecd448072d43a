#!/bin/bash
OUT=gpurun_out/align
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_scheme.py tests/test_gpu_sort.py tests/test_gpu_bootstrap.py tests/test_gpu_mehp24.py -m gpu -x -q > $OUT/pytest.log 2>&1
echo "parity exit $?" > $OUT/summary.txt
timeout 200 python tools/sort_once.py 1024 3 2>&1 | grep -E "N=|ops" | tail -n 2 >> $OUT/summary.txt
cat $OUT/summary.txt; tail -n 4 $OUT/pytest.log
