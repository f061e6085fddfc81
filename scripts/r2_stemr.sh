#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu --timeout 200 -p no:cacheprovider tests/test_gpu_ops.py -k "stem" -x 2>&1 | tail -8
timeout 600 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_refsize.py tests/test_gpu_e2e.py -x 2>&1 | tail -4
for m in yolov7-w6-custom yolov7-e6e; do
  for v in "" "YB_STEM_MMA_SYNC=1"; do
    env $v timeout 600 python bench.py --model $m --img 1280 --batch 8 --steps 20 --warmup 5 --no-extras 2>gpurun_out/p6.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$m $v value', round(d['value'],1), 'ms/step', round(d['ms_per_step'],4))" || tail -3 gpurun_out/p6.err
  done
  timeout 600 python scripts/profile_ops.py $m 8 1280 2>/dev/null | sed -n 2p
done
