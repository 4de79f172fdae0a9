#!/bin/bash
# ncu pass: launch list of the bench command + one full capture of the tcgen05 kernel and of the staged variance kernel.
TAG=${1:-prof}
OUT=gpurun_out/$TAG
mkdir -p $OUT
BCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $BCMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:psweep_tc_kernel -s 3 -c 1 -o $OUT/prof_tc $BCMD > $OUT/ncu_tc.log 2>&1
echo "ncu tc rc=$?"
VCMD="python tools/bench_kernels.py --reps 1 --only mvs_variance"
timeout 300 $VCMD > $OUT/plain_var.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mvs_variance_tile -s 2 -c 1 -o $OUT/prof_var $VCMD > $OUT/ncu_var.log 2>&1
echo "ncu var rc=$?"
ls -la $OUT
