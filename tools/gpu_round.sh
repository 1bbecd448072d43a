#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, ncu launch list of the bench's timed region, ncu --set full of one
# rotation's kernels.  Usage: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/gpu.txt 2>&1
nproc > $OUT/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/smoke.log 2>&1; echo "smoke exit $?" >> $OUT/smoke.log
tail -2 $OUT/smoke.log
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
cat $OUT/bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
cat $OUT/bench_ref.json | cut -c1-200
if [ "$2" != "noncu" ]; then
timeout 300 python bench.py --skip-sort --steps 2 --warmup 1 > $OUT/plain.log 2> $OUT/plain.err && {
  SKIP=$(grep -o 'timed region: [0-9]*' $OUT/plain.err | grep -o '[0-9]*$')
  echo "ncu skip $SKIP"
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c 400 --csv \
      --log-file $OUT/launches.csv python bench.py --skip-sort --steps 2 --warmup 1 > $OUT/ncu_launches.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -s $SKIP -c ${NCU_FULL_COUNT:-14} \
      -o $OUT/prof python bench.py --skip-sort --steps 2 --warmup 1 > $OUT/ncu_full.log 2>&1
  ls -la $OUT
}
fi
