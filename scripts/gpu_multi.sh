#!/bin/bash
# multi-GPU evidence: bit-identity of the sharded result + bench lines (run with gpurun --gpus N -- bash scripts/gpu_multi.sh N)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 scripts/check_dist_gpu.py yolov7-tiny 320 16 2>&1 | grep "dist check" | tee gpurun_out/dist_check_${N}gpu.txt
$TR --master-port 29512 scripts/check_dist_gpu.py yolov7-w6-custom 1280 16 2>&1 | grep "dist check" | tee -a gpurun_out/dist_check_${N}gpu.txt
$TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 3 --no-extras 2> gpurun_out/bench_${N}gpu.err | tail -1 > gpurun_out/bench_${N}gpu.json
$TR --master-port 29514 bench.py --gpus $N --steps 10 --warmup 3 --no-extras --model yolov7-w6-custom --img 1280 --global-batch 16 2> gpurun_out/bench_w6_${N}gpu.err | tail -1 > gpurun_out/bench_w6_${N}gpu.json
if [ "$N" = "8" ]; then
$TR --master-port 29515 bench.py --gpus $N --steps 10 --warmup 3 --no-extras --model yolov7-e6e --img 1280 --global-batch 64 2> gpurun_out/bench_e6e_${N}gpu.err | tail -1 > gpurun_out/bench_e6e_${N}gpu.json
fi
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/bench*_${N}gpu.json')):
    try:
        d=json.load(open(f)); print(f, d['metric'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['scaling'], d['config']['workload'])
    except Exception as e:
        print(f, 'parse failed', e)
PY
