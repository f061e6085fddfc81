#!/bin/bash
# usage: gpu_ncu_conv.sh <report-name> <spec>...   (plain run first, then one ncu --set full capture)
mkdir -p gpurun_out
name=$1; shift
YB_REPS=5 timeout 300 python scripts/bench_conv.py "$@" > gpurun_out/${name}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${name}_plain.log; exit 1; }
cat gpurun_out/${name}_plain.log
YB_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conv --launch-skip 0 \
  -o gpurun_out/${name} -f python scripts/bench_conv.py "$@" > gpurun_out/${name}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/${name}.ncu-rep
