set -x
O=gpurun_out/r02v; mkdir -p $O
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,launch__grid_size --clock-control none -k regex:psweep_ws -c 200 --csv --log-file $O/peer_local_ncu.csv python tools/peer_local_check.py --views 10 --peers 2 > $O/ncu.log 2>&1
python - <<PY
import csv,collections
rows=list(csv.reader(open("$O/peer_local_ncu.csv")))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]; kn=h.index('Kernel Name'); mn=h.index('Metric Name'); mv=h.index('Metric Value'); idc=h.index('ID')
d=collections.OrderedDict()
for r in rows[i+1:]:
    if len(r)>mv: d.setdefault(r[idc],{'k':r[kn][:30]})[r[mn]]=r[mv]
prev=None
for k,v in d.items():
    key=(v['k'],v.get('launch__grid_size'),v.get('smsp__inst_executed.sum'))
    if key!=prev: print(k,v)
    prev=key
PY
