#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_nms.py tests/test_gpu_fullsize.py -x 2>&1 | tail -5
timeout 300 python scripts/ml_time.py 2>&1 | tail -3
