#!/bin/bash
# ncu launch list + full capture of the correlation backward kernel (after a plain run of the same command exited 0).
TAG=${1:-bwdprof}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python tools/bench_kernels.py --reps 2 --only psweep_corr_bwd"
timeout 300 $CMD > $OUT/plain.log 2>&1 || { echo plain failed; tail -5 $OUT/plain.log; exit 1; }
cat $OUT/plain.log | cut -c1-300
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launches.log 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:psweep_corr_bwd_kernel -s 2 -c 1 -o $OUT/prof_bwd $CMD > $OUT/ncu_bwd.log 2>&1; echo "ncu rc=$?"
ls -la $OUT
