#!/bin/bash
mkdir -p gpurun_out
python bench.py --model ${MODEL:-yolov7-tiny} --batch 1 --steps 3 --warmup 3 --no-extras > gpurun_out/b1_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/b1_launches.csv python bench.py --model ${MODEL:-yolov7-tiny} --batch 1 --steps 3 --warmup 3 --no-extras > /dev/null 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(l for l in open('gpurun_out/b1_launches.csv') if not l.startswith('=='))]
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); ii=h.index('ID'); ui=h.index('Metric Unit')
d=[(int(r[ii]), r[ki].split('(')[0][:28], float(r[vi].replace(',',''))*({'ns':1e-3,'us':1,'ms':1e3}.get(r[ui],1))) for r in rows[1:]]
last=max(i for i,n,_ in d if 'sort_nms' in n); prev=max(i for i,n,_ in d if 'sort_nms' in n and i<last)
step=[x for x in d if prev<x[0]<=last]
print('launches', len(step), 'sum us', round(sum(x[2] for x in step),1))
from collections import defaultdict
a=defaultdict(lambda:[0,0.0])
for _,n,t in step: a[n][0]+=1; a[n][1]+=t
for n,(c,t) in sorted(a.items(), key=lambda kv:-kv[1][1]): print(f'{n:30s} {c:3d} {t:8.1f} us  avg {t/c:6.2f}')
PY
