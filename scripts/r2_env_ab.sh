#!/bin/bash
# same-box A/B of two environment settings with ONE library: VARS="A=1,B=2 C=3" (comma-separated assignments per arm; "-" = none)
mkdir -p gpurun_out
for r in $(seq ${ROUNDS:-2}); do for v in $VARS; do
  envs=$(echo $v | tr ',' ' '); [ "$v" = "-" ] && envs=""
  for cfg in ${CFGS:-yolov7:32:640}; do IFS=: read m b i <<< "$cfg"
    env $envs timeout 600 python bench.py --model $m --batch $b --img $i --steps ${STEPS:-60} --warmup 5 --no-extras 2>gpurun_out/q.err | tail -1 | python scripts/benchline.py "[$v]" $m B=$b $i || tail -5 gpurun_out/q.err
  done
done; done
