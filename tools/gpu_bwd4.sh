#!/bin/bash
TAG=${1:-bwd4}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_planesweep_bwd.py -q -x > $OUT/pytest_bwd.log 2>&1; echo "bwd pytest rc=$?"; tail -12 $OUT/pytest_bwd.log
for z in auto 1 2; do
  if [ $z = auto ]; then unset RMVD_BWD_ZSPLIT; else export RMVD_BWD_ZSPLIT=$z; fi
  timeout 200 python tools/bench_kernels.py --reps 10 --only psweep_corr_bwd 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('zsplit $z', round(d['ms_cold'],4), round(d['ms_warm'],4))"
done | tee $OUT/sweep.txt
unset RMVD_BWD_ZSPLIT
