#!/bin/bash
TAG=${1:-bwd3}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_models.py -q -x -s > $OUT/pytest_models.log 2>&1; echo "models rc=$?"; grep -E "model gradient errors|AssertionError|passed|failed" $OUT/pytest_models.log | cut -c1-900 | tail -6
for cfg in "32 320" "64 320" "64 512" "128 512" "128 800" "256 800" "256 1200"; do
  set -- $cfg
  RMVD_BWD_PLANES=$1 RMVD_BWD_RMAX=$2 timeout 200 python tools/bench_kernels.py --reps 10 --only psweep_corr_bwd 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('planes $1 rmax $2', round(d['ms_cold'],4), round(d['ms_warm'],4))"
done | tee $OUT/sweep.txt
CMD="python tools/bench_kernels.py --reps 2 --only psweep_corr_bwd"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:psweep_corr_bwd_tile -s 2 -c 1 -o $OUT/prof_bwd_tile $CMD > $OUT/ncu_bwd.log 2>&1; echo "ncu rc=$?"
