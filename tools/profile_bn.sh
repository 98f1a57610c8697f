#!/bin/bash
# per-launch time + DRAM bytes of the BatchNorm kernels of one ResNet training step (ncu, un-graphed)
set -u
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
  -k regex:'bn_' -s 1000 -c 340 --csv --log-file gpurun_out/bn_launches.csv python tools/profile_train.py 256 > gpurun_out/ncu_bn.log 2>&1; echo rc=$?
python - <<'PY'
import csv
rows={}
for r in csv.DictReader(l for l in open('gpurun_out/bn_launches.csv') if l.startswith('"')):
    d=rows.setdefault(int(r['ID']),{'k':r['Kernel Name'].split('(')[0][:40]})
    d[r['Metric Name']]=float(r['Metric Value'].replace(',',''))
from collections import defaultdict
agg=defaultdict(list)
for d in rows.values():
    t=d.get('gpu__time_duration.sum',0)/1e3; by=(d.get('dram__bytes_read.sum',0)+d.get('dram__bytes_write.sum',0))/1e6
    agg[d['k']].append((t,by))
for k,v in agg.items():
    big=[(t,b) for t,b in v if b>150]
    print(f"{k:42s} n={len(v):4d} total {sum(t for t,_ in v):8.1f} us  {sum(b for _,b in v)/1e3:7.2f} GB -> {sum(b for _,b in v)/max(sum(t for t,_ in v),1e-9)/1e3:6.2f} TB/s ; launches > 150 MB: n={len(big)} {sum(b for _,b in big)/max(sum(t for t,_ in big),1e-9)/1e3:6.2f} TB/s")
PY
