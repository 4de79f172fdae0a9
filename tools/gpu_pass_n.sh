#!/bin/bash
# bench.py on N GPUs of one box, launched as the driver launches it.   gpurun --gpus N --timeout 900 -- 'bash tools/gpu_pass_n.sh N tag'
N=$1; TAG=${2:-pass}; O=gpurun_out/$TAG; mkdir -p $O
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?"
tail -3 $O/bench_n$N.err
python - <<PY
import json
d=json.load(open("$O/bench_n$N.json"))
print({k:d[k] for k in ('value','ms_per_step','n_gpus')}, 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
sl=d.get('slab',{})
for r in sl.get('planesweep',[]): print(r['views'], round(r['ms_1gpu'],4), {k:(r[k]['ms_total'] and round(r[k]['ms_total'],4), round(r[k]['speedup'],3), r[k]['bit_exact'], r[k]['mode'][:30]) for k in ('planes','tiles','planes_copy','planes_multicast','tiles_multicast')})
print(sl.get('regress'), sl.get('error'))
print([ (c['S'],c['V'],round(c['voxels_per_s']/1e9,1)) for c in d.get('sweep',{}).get('cells',[])][:20])
PY
