set -x
O=gpurun_out/r02r; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_mvs.py tests/test_gpu_models.py -x -q > $O/pytest_mvs.log 2>&1; echo "pytest rc=$?" >> $O/pytest_mvs.log
tail -15 $O/pytest_mvs.log
for m in 0 1 2 3; do RMVD_MVS_VARIANCE_KERNEL=$m timeout 300 python tools/bench_kernels.py --only mvs_variance > $O/var_mode$m.jsonl 2>&1; cat $O/var_mode$m.jsonl | cut -c1-300; done
for s in 2 8; do RMVD_MVS_VARIANCE_SPLIT=$s timeout 300 python tools/bench_kernels.py --only mvs_variance > $O/var_split$s.jsonl 2>&1; cat $O/var_split$s.jsonl | cut -c1-300; done
ncu --set full --clock-control none --import-source on -k regex:mvs_variance_cl -c 1 -o $O/var_full python tools/bench_kernels.py --only mvs_variance --reps 1 > $O/ncu_var.log 2>&1
