#!/bin/bash
# quick GPU iteration: selected suites, then short bench lines without extras.  TESTS="..." CFGS="model:batch:img ..."
mkdir -p gpurun_out
TESTS=${TESTS:-"tests/test_gpu_ops.py tests/test_gpu_e2e.py tests/test_gpu_refsize.py"}
timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider $TESTS -x 2>&1 | tail -6
for cfg in ${CFGS:-yolov7-tiny:1:640 yolov7:1:640 yolov7:32:640}; do IFS=: read m b i <<< "$cfg"
  timeout 600 python bench.py --model $m --batch $b --img $i --steps 30 --warmup 5 --no-extras 2>gpurun_out/q.err | tail -1 | python scripts/benchline.py $m B=$b $i || tail -5 gpurun_out/q.err
done
if [ -n "$PROFILE" ]; then timeout 600 python scripts/profile_ops.py yolov7 32 640 > gpurun_out/ops.tsv 2>gpurun_out/ops.err; tail -1 gpurun_out/ops.tsv; fi
