#!/bin/bash
# Round-2 standard GPU check: the GPU suite exactly as the driver runs it, smoke, the bench line with extras.
mkdir -p gpurun_out; rm -f gpurun_out/parity_refsize.jsonl gpurun_out/summary.txt
timeout 1500 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "pytest -m gpu exit $? $(tail -n 1 gpurun_out/pytest_gpu.log)" | tee -a gpurun_out/summary.txt
grep -h "Error\|FAILED\|assert" gpurun_out/pytest_gpu.log | head -20
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/summary.txt; tail -n 2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" | tee -a gpurun_out/summary.txt
tail -c 1500 gpurun_out/bench.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'], 1), 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), 'clocks', d['clocks'])
r = d.get('roofline', {})
print({k: r.get(k) for k in ('kernel', 'bound', 'achieved', 'frac', 'ms_per_step', 'frac_of_layer_bound', 'trunk_graph_ms', 'decode_nms_ms', 'eager_per_op_sum_ms', 'conv_tflops', 'conv_frac_of_sustained_bf16', 'step_frac_of_layer_bounds')})
for k in d.get('roofline_kernels', []):
    print(' ', k['kernel'], k['launches'], 'ms', k['ms'], k['bound'], k['achieved'], k['unit'], 'frac', k['frac'], 'of-layer-bound', k['frac_of_layer_bound'], '| in graph ms', k.get('in_graph_ms'), 'of-layer-bound', k.get('in_graph_frac_of_layer_bound'))
ns = d.get('nms_stress') or {}
print('nms_stress', {k: ns.get(k) for k in ('us_per_image', 'frac_of_hbm_peak', 'cpu_us_per_image', 'cpu_us_per_image_1thread', 'cpu_threads', 'gpu_vs_cpu', 'parity_on_sample', 'error')})
print('cpu_baseline', d.get('cpu_baseline'))
PY
cat gpurun_out/parity_refsize.jsonl 2>/dev/null | cut -c1-260
