#!/bin/bash
# timing (and optionally a full-section capture) of the decode/filter + sort/NMS kernels inside a bench step
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/nn_plain.json 2> gpurun_out/nn_plain.err || { echo plain failed; exit 1; }
ncu ${NCU_SET:---metrics gpu__time_duration.sum} --clock-control none -k regex:'k_filter|k_sort' --import-source on \
  -o gpurun_out/$1 -f python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/$1_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/$1.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); hdr=rows[0]
for r in rows[2:]: print(r[hdr.index('Kernel Name')][:40], r[hdr.index('Grid Size')], r[hdr.index('gpu__time_duration.sum')])
"
