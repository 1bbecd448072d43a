#!/bin/bash
# ncu captures of round 2 (one gpurun call): --set full of one rotation + one EvalMult, of the kernels outside the rotation,
# and the launch list of one DirectSort N = 256.  Every profiled command first runs plain and must exit 0.
OUT=gpurun_out/prof2
mkdir -p $OUT
python tools/rot_once.py 1 > $OUT/rot_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -o $OUT/rot python tools/rot_once.py 1 > $OUT/rot_ncu.log 2>&1
python tools/prof_misc.py > $OUT/misc_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:'lc_tc|lc_build|ew_rescale|ew_tensor|enc_|ks_dh|ew_mul_pt|ew_binary|ew_mul_scalar|ew_automorphism' -c 60 \
    -o $OUT/misc python tools/prof_misc.py > $OUT/misc_ncu.log 2>&1
python tools/sort_once.py 256 2 > $OUT/sort256_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $OUT/sort256_launches.csv \
    python tools/sort_once.py 256 2 > $OUT/sort256_ncu.log 2>&1
# summaries here (the reports are too large to travel back: 64 MiB limit), then drop the reports
python tools/ncu_raw.py $OUT/rot.ncu-rep > $OUT/ncu_full_rot.txt 2>&1
python tools/ncu_raw.py $OUT/misc.ncu-rep > $OUT/ncu_full_misc.txt 2>&1
ncu -i $OUT/rot.ncu-rep --page source --csv --kernel-name regex:ntt_fwd_pass_c 2>/dev/null | gzip > $OUT/source_ntt_fwd.csv.gz
ncu -i $OUT/rot.ncu-rep --page source --csv --kernel-name regex:bc_convert 2>/dev/null | gzip > $OUT/source_bc_convert.csv.gz
python tools/ncu_launch_summary.py $OUT/sort256_launches.csv > $OUT/sort256_kernels.txt 2>&1
rm -f $OUT/rot.ncu-rep $OUT/misc.ncu-rep
tail -n 2 $OUT/*_plain.log
ls -la $OUT
