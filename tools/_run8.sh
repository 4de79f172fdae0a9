set -x
O=gpurun_out/r02z; mkdir -p $O
RMVD_MVS_VARIANCE_KERNEL=4 timeout 180 python -m pytest tests/test_gpu_mvs.py -x -q -k "variance or cfg3" > $O/pytest_mvs_ws.log 2>&1; echo "pytest rc=$?" >> $O/pytest_mvs_ws.log
tail -5 $O/pytest_mvs_ws.log
for m in 4 5 0; do RMVD_MVS_VARIANCE_KERNEL=$m timeout 120 python tools/bench_kernels.py --only mvs_variance > $O/var_mode$m.jsonl 2>&1; cat $O/var_mode$m.jsonl | cut -c1-200; done
for s in 1 2 4; do RMVD_MVS_VARIANCE_KERNEL=4 RMVD_MVS_VARIANCE_SPLIT=$s timeout 120 python tools/bench_kernels.py --only mvs_variance > $O/var_ws_split$s.jsonl 2>&1; cat $O/var_ws_split$s.jsonl | cut -c1-200; done
RMVD_MVS_VARIANCE_KERNEL=4 timeout 300 ncu --set full --clock-control none --import-source on -k regex:mvs_variance_ws -c 1 -o $O/var_ws_full python tools/bench_kernels.py --only mvs_variance --reps 1 > $O/ncu_var.log 2>&1
