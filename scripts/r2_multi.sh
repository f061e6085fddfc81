#!/bin/bash
# N-GPU bench lines (weak scaling headline + configs[2] strong scaling), run under torchrun like the driver does
N=${N:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-extras > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "weak rc=$?"
tail -c 600 gpurun_out/bench_${N}gpu.err
python - <<PY
import json
d = json.loads(open('gpurun_out/bench_${N}gpu.json').read().strip().splitlines()[-1])
print('N', d['n_gpus'], 'value', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'gather', d['config']['gather'], 'gather_check', d.get('gather_check'))
PY
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --model yolov7-w6-custom --img 1280 --global-batch 16 > gpurun_out/bench_w6_${N}gpu.json 2> gpurun_out/bench_w6_${N}gpu.err; echo "w6 rc=$?"
tail -c 600 gpurun_out/bench_w6_${N}gpu.err
python - <<PY
import json
d = json.loads(open('gpurun_out/bench_w6_${N}gpu.json').read().strip().splitlines()[-1])
print('w6 N', d['n_gpus'], 'value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'gather_check', d.get('gather_check'))
r = d.get('roofline', {})
print({k: r.get(k) for k in ('kernel', 'frac', 'frac_of_layer_bound', 'conv_tflops', 'conv_frac_of_sustained_bf16', 'step_frac_of_layer_bounds')})
PY
timeout 300 $TR bench.py --impl reference --gpus $N --steps 1 --warmup 0 | tail -c 400
