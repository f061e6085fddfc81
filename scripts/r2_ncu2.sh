#!/bin/bash
# ncu --set full captures of the round-2 kernel variants (CTA pairs, P6 stem), each after the same command has passed
mkdir -p gpurun_out
export YB_REPS=2
A='python scripts/bench_conv.py 1024,1024,1,1,40'
B='python scripts/bench_conv.py 256,256,3,1,80'
C='python bench.py --model yolov7-w6-custom --img 1280 --batch 8 --steps 1 --warmup 3 --no-extras --streams 1'
$A > gpurun_out/plain2_a.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_conv_tc -s 3 -c 1 -o gpurun_out/r2_conv_tc_pair_1024x1024_40 -f $A > gpurun_out/ncu2_a.log 2>&1
echo "a rc=$?"
$B > gpurun_out/plain2_b.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_conv3_tc -s 3 -c 1 -o gpurun_out/r2_conv3_tc_pair_256x256_80 -f $B > gpurun_out/ncu2_b.log 2>&1
echo "b rc=$?"
$C > gpurun_out/plain2_c.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_stemr_tc -s 3 -c 1 -o gpurun_out/r2_stemr_tc -f $C > gpurun_out/ncu2_c.log 2>&1
echo "c rc=$?"
ls -la gpurun_out/r2_*pair*.ncu-rep gpurun_out/r2_stemr_tc.ncu-rep
