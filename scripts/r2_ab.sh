#!/bin/bash
# same-box A/B of two builds of the library: scripts/ab/base.so and scripts/ab/new.so alternate (boxes differ by 2-3 %
# under the power cap, so only runs of one gpurun call compare).  CFGS="model:batch:img ..."  ROUNDS=2
mkdir -p gpurun_out
for r in $(seq ${ROUNDS:-2}); do for v in base new; do
  cp scripts/ab/$v.so yolo_series_b200/libyolo_b200.so
  for cfg in ${CFGS:-yolov7:32:640}; do IFS=: read m b i <<< "$cfg"
    timeout 600 python bench.py --model $m --batch $b --img $i --steps 40 --warmup 5 --no-extras 2>gpurun_out/q.err | tail -1 | python scripts/benchline.py $v $m B=$b $i || tail -5 gpurun_out/q.err
  done
done; done
cp scripts/ab/new.so yolo_series_b200/libyolo_b200.so
