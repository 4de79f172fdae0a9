#!/bin/bash
# Final state check: all GPU tests, smoke, bench (both arms), every kernel timing.
TAG=${1:-final}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log; tail -4 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log; tail -3 $OUT/smoke.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cut -c1-900 $OUT/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"
timeout 600 python tools/bench_kernels.py --reps 20 > $OUT/kernels.jsonl 2> $OUT/kernels.err; echo "kernels rc=$?"; cut -c1-160 $OUT/kernels.jsonl; tail -3 $OUT/kernels.err
