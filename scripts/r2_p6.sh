#!/bin/bash
# P6 configs on one GPU (per-GPU shards of configs[2]/[3]): pair rule on/off, per-op table
mkdir -p gpurun_out
for m in yolov7-w6-custom yolov7-e6e; do
  for pr in 0 -1; do
    YB_PAIR=$pr timeout 600 python bench.py --model $m --img 1280 --batch 8 --steps 20 --warmup 5 --no-extras 2>gpurun_out/p6.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$m YB_PAIR=$pr value', round(d['value'],1), 'ms/step', round(d['ms_per_step'],4))" || tail -3 gpurun_out/p6.err
  done
  timeout 600 python scripts/profile_ops.py $m 8 1280 > gpurun_out/ops_$m.tsv 2>gpurun_out/p6.err || tail -3 gpurun_out/p6.err
  sort -t$'\t' -k9 -n -r gpurun_out/ops_$m.tsv | head -12; tail -1 gpurun_out/ops_$m.tsv
done
