#!/bin/bash
# launch list of a bench step (+ summary) and one full bench line, for profiles/
mkdir -p gpurun_out
bash scripts/gpu_launch_list.sh
python scripts/summarise_launches.py gpurun_out/launches.csv gpurun_out/launch_summary.json > /dev/null
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench.json'))
print(round(d['value'], 1), round(d['e2e']['value'], 1), d['roofline']['achieved'], d['roofline']['frac'], d['nms_stress']['us_per_image'],
      d['cpu_baseline']['value'], d['clocks'])
s = json.load(open('gpurun_out/launch_summary.json'))
print(s['launches_per_step'], s['sum_kernel_us'])
for k in s['kernels']:
    print('| `%s` | %d | %.1f | %.1f %% | %d |' % (k['kernel'][:28], k['launches'], k['us'], 100 * k['share'], round(k['dram_bytes'] / 1e6)))
PY
