set -x
O=gpurun_out/r03a; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
for m in 0 1 2 9; do RMVD_MVS_VARIANCE_KERNEL=$m timeout 120 python tools/bench_kernels.py --only mvs_variance > $O/var_mode$m.jsonl 2>&1; cat $O/var_mode$m.jsonl | cut -c1-200; done
timeout 300 ncu --set full --clock-control none --import-source on -k regex:mvs_variance_ws -c 1 -o $O/var_ws_full python tools/bench_kernels.py --only mvs_variance --reps 1 > $O/ncu_var.log 2>&1
timeout 600 python tools/bench_kernels.py > $O/kernels.jsonl 2> $O/kernels.err
