#!/bin/bash
# One gpurun call: parity tests, smoke, bench (both arms), per-kernel timings, then the ncu launch list and full captures of the
# two dominant kernels.  Usage (from the repo root on the GPU box): bash tools/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt; lscpu | grep -E 'Model name|^CPU\(s\)' >> $OUT/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
timeout 200 python tools/tc_soak.py 200 > $OUT/soak.log 2>&1; echo "soak rc=$?"; tail -1 $OUT/soak.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"; cat $OUT/bench_ref.json
timeout 600 python tools/bench_kernels.py --reps 20 > $OUT/kernels.jsonl 2> $OUT/kernels.err; echo "kernels rc=$?"
timeout 100 python tools/tc_stats8.py 8 > $OUT/tc_stats8.log 2>&1; tail -4 $OUT/tc_stats8.log
BCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $BCMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $BCMD > $OUT/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:psweep_tc_kernel -s 3 -c 1 -o $OUT/prof_tc $BCMD > $OUT/ncu_tc.log 2>&1
echo "ncu tc rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:nchw_to_split -s 3 -c 1 -o $OUT/prof_repack $BCMD > $OUT/ncu_repack.log 2>&1
VCMD="python tools/bench_kernels.py --reps 1 --only mvs_variance"
timeout 300 $VCMD > $OUT/plain_var.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mvs_variance_tile -s 2 -c 1 -o $OUT/prof_var $VCMD > $OUT/ncu_var.log 2>&1
echo "ncu var rc=$?"
WCMD="python tools/bench_kernels.py --reps 1 --only psweep_corr_bwd_cfg2_kitti_1frame"
timeout 300 $WCMD > $OUT/plain_bwd.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:psweep_corr_bwd_tile -s 1 -c 1 -o $OUT/prof_bwd $WCMD > $OUT/ncu_bwd.log 2>&1
echo "ncu bwd rc=$?"
ls -la $OUT
