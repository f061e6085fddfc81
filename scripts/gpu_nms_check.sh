#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_nms.py tests/test_gpu_fullsize.py tests/test_gpu_e2e.py -x 2>&1 | tail -3
YB_NMS_SORT_THREADS=512 timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_nms.py -x 2>&1 | tail -2
python scripts/nms_stress_only.py 2>&1 | tail -1
bash scripts/gpu_ncu_nms.sh nn_now 2>&1 | tail -2
