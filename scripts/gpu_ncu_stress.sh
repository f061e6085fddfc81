#!/bin/bash
mkdir -p gpurun_out
python scripts/nms_stress_only.py > gpurun_out/stress_plain.json 2> gpurun_out/stress_plain.err || { tail -5 gpurun_out/stress_plain.err; exit 1; }
cat gpurun_out/stress_plain.json
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_filter|k_sort|k_' --csv --log-file gpurun_out/stress_launches.csv python scripts/nms_stress_only.py > gpurun_out/stress_ncu.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/stress_launches.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); mi=hdr.index('Metric Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID'); gi=hdr.index('Grid Size')
d={}
for r in rows[1:]:
    d.setdefault((r[ii],r[ki][:40],r[gi]),{})[r[mi]]=r[vi]
for k,v in list(d.items())[-8:]:
    print(k, v)
PY
