set -x
N=$1; O=gpurun_out/r02w; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?"
tail -5 $O/bench_n$N.err
python - <<PY
import json
d=json.load(open("$O/bench_n$N.json"))
print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'])
print(json.dumps(d.get('slab'),indent=1)[:5000])
PY
