#!/bin/bash
TAG=${1:-bwd2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_planesweep_bwd.py -q -x > $OUT/pytest_bwd.log 2>&1; echo "bwd pytest rc=$?"; tail -25 $OUT/pytest_bwd.log
timeout 300 python -m pytest tests/test_gpu_models.py -q -x > $OUT/pytest_models.log 2>&1; echo "models rc=$?"; grep -E "AssertionError|assert |passed|failed" $OUT/pytest_models.log | tail -8
timeout 300 python tools/bench_kernels.py --reps 10 --only psweep_corr_bwd > $OUT/kernels_bwd.jsonl 2> $OUT/kernels_bwd.err; echo "kernels rc=$?"; cut -c1-400 $OUT/kernels_bwd.jsonl; tail -3 $OUT/kernels_bwd.err
RMVD_PSWEEP_BWD_KERNEL=direct timeout 300 python tools/bench_kernels.py --reps 10 --only psweep_corr_bwd > $OUT/kernels_bwd_direct.jsonl 2>&1; cut -c1-200 $OUT/kernels_bwd_direct.jsonl
