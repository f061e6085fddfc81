#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest -q -m gpu --timeout 120 -p no:cacheprovider tests/test_gpu_ops.py -k "cta_pairs" -x 2>&1 | tail -25
