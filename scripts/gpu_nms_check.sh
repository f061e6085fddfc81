#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_nms.py tests/test_gpu_fullsize.py tests/test_gpu_tta.py -x 2>&1 | tail -5
YB_NMS_SORT_THREADS=512 timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_nms.py tests/test_gpu_e2e.py -x 2>&1 | tail -5
python scripts/nms_stress_only.py 2>&1 | tail -1
python bench.py --steps 20 --warmup 3 --no-extras 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bench value', d['value'], 'e2e', d['e2e']['value'])"
