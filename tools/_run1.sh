set -x
O=gpurun_out/r02n; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py > $O/bench.json 2> $O/bench.err
python tools/bench_kernels.py > $O/kernels.jsonl 2> $O/kernels.err
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/b_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:psweep_ws_kernel -s 2 -c 1 -o $O/ws_full python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:nchw_to_grouped -s 2 -c 1 -o $O/grouped_full python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/ncu_full2.log 2>&1
tail -3 $O/pytest_gpu.log; cat $O/smoke.log | tail -2; cat $O/bench.json | cut -c1-600
