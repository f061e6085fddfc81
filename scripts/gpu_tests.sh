#!/bin/bash
# Runs the GPU suites one process per file (a hung kernel only loses that file), logs to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
PT="python -m pytest -q -m gpu --timeout 300 --timeout-method=thread -p no:cacheprovider"
for f in test_gpu_nms test_gpu_ops test_gpu_e2e test_gpu_pre test_gpu_tta test_gpu_eval test_gpu_fullsize; do
  timeout 900 $PT tests/$f.py -s > gpurun_out/$f.log 2>&1
  echo "$f exit $?" | tee -a gpurun_out/summary.txt
  tail -n 25 gpurun_out/$f.log
done
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/summary.txt; tail -n 5 gpurun_out/smoke.log
