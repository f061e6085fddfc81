#!/bin/bash
# 8-GPU lines: weak-scaling headline, configs[2] (w6-custom 1280 B=16 strong), configs[3] (e6e 1280 B=64 = 8/GPU)
N=${N:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() {  # name, bench args
  name=$1; shift
  timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_${N}gpu.json 2> gpurun_out/bench_${name}_${N}gpu.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_${name}_${N}gpu.json').read().strip().splitlines()[-1])
    r = d.get('roofline', {})
    print('$name N', d['n_gpus'], 'value', round(d['value'], 1), 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value'], 1), 'gather_check', d.get('gather_check'),
          {k: r.get(k) for k in ('kernel', 'frac', 'conv_tflops', 'conv_frac_of_sustained_bf16', 'step_frac_of_layer_bounds')})
except Exception as ex:
    print('$name: no line', ex)
PY
  tail -c 300 gpurun_out/bench_${name}_${N}gpu.err | grep -v "OMP_NUM_THREADS\|\*\*\*\*" | tail -3
}
run v7 --no-extras
run w6 --model yolov7-w6-custom --img 1280 --global-batch 16
run e6e --model yolov7-e6e --img 1280 --global-batch 64
