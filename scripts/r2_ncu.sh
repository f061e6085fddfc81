#!/bin/bash
# ncu --set full captures of the round-2 kernels (one launch each), each after the same command has passed without ncu
mkdir -p gpurun_out
export YB_REPS=2
A='python scripts/bench_conv.py 128,128,1,1,160'
B='python scripts/bench_conv.py 64,64,3,1,320'
C='python bench.py --steps 1 --warmup 3 --no-extras'
$A > gpurun_out/plain_a.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_conv_tc -s 3 -c 1 -o gpurun_out/r2_conv_tc_128x128_160 -f $A > gpurun_out/ncu_a.log 2>&1
echo "a rc=$?"
$B > gpurun_out/plain_b.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_conv3_tc -s 3 -c 1 -o gpurun_out/r2_conv3_tc_64x64_320 -f $B > gpurun_out/ncu_b.log 2>&1
echo "b rc=$?"
$C > gpurun_out/plain_c.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_stem_tc -s 3 -c 1 -o gpurun_out/r2_stem_tc -f $C > gpurun_out/ncu_c.log 2>&1
echo "c rc=$?"
ls -la gpurun_out/*.ncu-rep; tail -3 gpurun_out/ncu_a.log gpurun_out/ncu_c.log
