#!/bin/bash
# One gpurun call: parity tests, smoke, bench, then the ncu launch list and one full capture of the top kernel.
# Usage (from the repo root on the GPU box): bash tools/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt; lscpu | grep -E 'Model name|^CPU\(s\)' >> $OUT/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"; cat $OUT/bench_ref.json
BCMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $BCMD > $OUT/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $BCMD > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $BCMD > $OUT/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:psweep_tc -s 3 -c 2 -o $OUT/prof_tc $BCMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
