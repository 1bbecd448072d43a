#!/bin/bash
# default (auto) selection of the base-conversion kernels: parity tests, kbench, sorts
OUT=gpurun_out/bc5auto
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_kernels_limb.py tests/test_gpu_primitives.py tests/test_gpu_scheme.py -m gpu -x -q -k "not kway" > $OUT/pytest.log 2>&1
echo "parity exit $?" > $OUT/summary.txt
timeout 300 python tools/kbench.py auto >> $OUT/summary.txt 2>> $OUT/err.txt
timeout 300 python tools/sort_once.py 256 3 2>&1 | tail -n 2 >> $OUT/summary.txt
timeout 300 python tools/sort_once.py 1024 3 2>&1 | tail -n 2 >> $OUT/summary.txt
cat $OUT/summary.txt; tail -n 3 $OUT/pytest.log
