#!/bin/bash
timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_ops.py tests/test_gpu_e2e.py tests/test_gpu_nms.py -x 2>&1 | tail -4
for cfg in "yolov7-tiny 1" "yolov7 32"; do set -- $cfg; python bench.py --model $1 --batch $2 --steps 50 --warmup 5 --no-extras 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:32], '| value', round(d['value'],1), 'ms/step', round(d['ms_per_step'],3), '| e2e', round(d['e2e']['value'],1))"; done
python scripts/profile_ops.py yolov7 32 640 2>/dev/null | grep "k4\|total"
