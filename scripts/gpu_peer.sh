#!/bin/bash
# in-kernel peer gather vs NCCL all-gather on N GPUs: correctness, then bench lines of both
N=${1:-2}
mkdir -p gpurun_out
TR="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29521 scripts/check_dist_gpu.py yolov7-tiny 320 16 2>&1 | grep -E "dist check|Error|error" | head -5 | tee gpurun_out/dist_check_${N}gpu.txt
$TR --master-port 29522 bench.py --gpus $N --steps 30 --warmup 5 --no-extras 2> gpurun_out/bench_peer_${N}.err | tail -1 > gpurun_out/bench_peer_${N}gpu.json
YB_GATHER=nccl $TR --master-port 29523 bench.py --gpus $N --steps 30 --warmup 5 --no-extras 2> gpurun_out/bench_nccl_${N}.err | tail -1 > gpurun_out/bench_nccl_${N}gpu.json
tail -3 gpurun_out/bench_peer_${N}.err
python - <<PY
import json
for k in ('peer','nccl'):
    try:
        d=json.load(open(f'gpurun_out/bench_{k}_${N}gpu.json')); print(k, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'serial', round(d['e2e']['serial_value'],1), d['config'].get('gather'))
    except Exception as e: print(k, 'failed', e)
PY
