#!/bin/bash
# One gpurun call for the correlation backward: its parity tests first, then the whole GPU suite, kernel timings, bench.
TAG=${1:-bwd}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_planesweep_bwd.py -q -x > $OUT/pytest_bwd.log 2>&1; echo "bwd pytest rc=$?"; tail -15 $OUT/pytest_bwd.log
timeout 300 python tools/bench_kernels.py --reps 10 --only psweep_corr_bwd > $OUT/kernels_bwd.jsonl 2> $OUT/kernels_bwd.err; echo "kernels rc=$?"; cut -c1-400 $OUT/kernels_bwd.jsonl; tail -3 $OUT/kernels_bwd.err
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cut -c1-1200 $OUT/bench.json
