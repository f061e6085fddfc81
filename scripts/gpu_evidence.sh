#!/bin/bash
# Refreshes the judged evidence for the current code: launch list of a bench step (+ summary), ncu --set full of the
# NMS kernels inside a bench step and on the stress config, and a full bench line.  Each ncu pass runs only after the
# same command has exited 0 without ncu.
mkdir -p gpurun_out
bash scripts/gpu_launch_list.sh
python scripts/summarise_launches.py gpurun_out/launches.csv gpurun_out/launch_summary.json && head -c 300 gpurun_out/launch_summary.json; echo
NCU_SET="--set full" bash scripts/gpu_ncu_nms.sh nms_step_full | tail -3
python scripts/ncu_summary.py gpurun_out/nms_step_full.ncu-rep 2 3 > gpurun_out/nms_step_summary.txt
bash scripts/gpu_ncu_stress_full.sh > /dev/null 2>&1; head -8 gpurun_out/stress_full_summary.txt
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench.json'))
print({k: d[k] for k in ('value', 'ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', d['roofline']['achieved'], d['roofline']['frac'], 'nms', d['nms_stress']['us_per_image'], 'cpu', d['cpu_baseline'])
PY
