#!/bin/bash
# ncu --set full of the stand-alone NMS kernels on BASELINE configs[4] (after the plain run exited 0)
mkdir -p gpurun_out
python scripts/nms_stress_only.py > gpurun_out/stress_plain.json 2> gpurun_out/stress_plain.err || { tail -5 gpurun_out/stress_plain.err; exit 1; }
cat gpurun_out/stress_plain.json
ncu --set full --clock-control none --import-source on -k regex:'k_filter_bulk|k_sort_nms' -s 4 -c 2 -o gpurun_out/stress_full -f \
  python scripts/nms_stress_only.py > gpurun_out/stress_full_ncu.log 2>&1
echo "ncu rc=$?"
python scripts/ncu_summary.py gpurun_out/stress_full.ncu-rep > gpurun_out/stress_full_summary.txt
cat gpurun_out/stress_full_summary.txt
