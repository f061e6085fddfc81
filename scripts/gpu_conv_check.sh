#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu tests/test_gpu_ops.py --timeout 200 --timeout-method=thread -p no:cacheprovider > gpurun_out/ops_test.log 2>&1
echo "ops tests rc=$?"; tail -5 gpurun_out/ops_test.log
timeout 300 python scripts/bench_conv.py "$@" 2>&1 | tee gpurun_out/conv_bench.log | tail -14
