#!/bin/bash
# 1-GPU bench lines WITH extras (per-kernel rooflines) for the P6 configs, and the small-batch latencies
mkdir -p gpurun_out
timeout 600 python bench.py --model yolov7-w6-custom --img 1280 --batch 16 --steps 20 --warmup 5 > gpurun_out/bench_w6_b16_1gpu.json 2>gpurun_out/p6.err; echo "w6 b16 rc=$?"
timeout 600 python bench.py --model yolov7-e6e --img 1280 --batch 8 --steps 20 --warmup 5 > gpurun_out/bench_e6e_b8_1gpu.json 2>>gpurun_out/p6.err; echo "e6e b8 rc=$?"
timeout 600 python bench.py --model yolov7-tiny --img 640 --batch 32 --steps 20 --warmup 5 > gpurun_out/bench_tiny_b32_1gpu.json 2>>gpurun_out/p6.err; echo "tiny b32 rc=$?"
python - <<'PY'
import json
for f in ('bench_w6_b16_1gpu', 'bench_e6e_b8_1gpu', 'bench_tiny_b32_1gpu'):
    d = json.loads(open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1])
    r = d['roofline']
    print(f, 'value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1),
          {k: r.get(k) for k in ('kernel', 'frac', 'conv_tflops', 'conv_frac_of_sustained_bf16', 'step_frac_of_layer_bounds')})
    for k in d['roofline_kernels'][:6]:
        print('   ', k['kernel'], k['launches'], 'ms', k['ms'], k['bound'], 'frac', k['frac'], 'of-layer-bound', k['frac_of_layer_bound'])
PY
