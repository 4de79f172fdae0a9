#!/bin/bash
# One measurement pass on a B200 box (run through gpurun): GPU tests, smoke, both bench arms, per-kernel timings, the ncu launch list
# and one full capture of the headline kernel.  Everything lands in gpurun_out/<tag>/; copy what is to be judged into profiles/.
#   gpurun --timeout 1500 -- 'bash tools/gpu_pass.sh r03f'
TAG=${1:-pass}; O=gpurun_out/$TAG; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py > $O/bench.json 2> $O/bench.err
python tools/bench_kernels.py > $O/kernels.jsonl 2> $O/kernels.err
python tools/profile_model_run.py > $O/model_run.json 2> $O/model_run.err
python tools/pcie_probe.py > $O/pcie.json 2> $O/pcie.err
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/b_small.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:psweep_ws_kernel -s 2 -c 1 -o $O/ws_full python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:nchw_to_keytiles -s 2 -c 1 -o $O/keytiles_full python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/ncu_full3.log 2>&1
tail -3 $O/pytest_gpu.log; tail -1 $O/smoke.log; cut -c1-400 $O/bench.json
