#!/bin/bash
N=${1:-4}
TR="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for g in peer nccl peer; do
i=$((i+1))
YB_GATHER=$g $TR --master-port $((29540+i)) bench.py --gpus $N --steps 30 --warmup 5 --no-extras --model yolov7-w6-custom --img 1280 --global-batch 16 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$g', 'value', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'serial', round(d['e2e']['serial_value'],1))"
done
