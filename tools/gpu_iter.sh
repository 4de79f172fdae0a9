#!/bin/bash
# Quick iteration call: stats run of the tcgen05 kernel, GPU parity tests, per-kernel timings, short bench.
TAG=${1:-iter}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python tools/tc_debug.py > $OUT/tc_debug.log 2>&1; echo "tc_debug rc=$?"; tail -12 $OUT/tc_debug.log
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
timeout 600 python tools/bench_kernels.py --reps 10 > $OUT/kernels.jsonl 2> $OUT/kernels.err; echo "kernels rc=$?"; cut -c1-200 $OUT/kernels.jsonl
RMVD_MVS_VARIANCE_DIRECT=1 timeout 300 python tools/bench_kernels.py --reps 5 --only mvs_variance > $OUT/kernels_direct.jsonl 2>&1; cut -c1-200 $OUT/kernels_direct.jsonl
timeout 600 python bench.py --steps 20 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cut -c1-1500 $OUT/bench.json
