#!/bin/bash
# tests + per-op profile + bench; logs into gpurun_out/ (run from the repo root on the GPU box)
mkdir -p gpurun_out
bash scripts/gpu_tests.sh > gpurun_out/tests_all.log 2>&1
cat gpurun_out/summary.txt
python scripts/profile_ops.py yolov7 32 640 > gpurun_out/ops.tsv 2> gpurun_out/ops.err
tail -1 gpurun_out/ops.tsv
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"
tail -c 600 gpurun_out/bench.err
python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/bench.json'))
    r = d.get('roofline', {})
    print({k: d[k] for k in ('value', 'ms_per_step')}, 'e2e', d['e2e']['value'], 'roof', r.get('achieved'), r.get('frac'),
          'nms', d.get('nms_stress'), 'cpu', d.get('cpu_baseline', {}).get('value'))
except Exception as e:
    print('bench parse failed', e)
PY
