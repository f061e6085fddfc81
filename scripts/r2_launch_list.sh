#!/bin/bash
# ncu launch list of bench.py's step: per-launch duration and DRAM bytes, taken only after the same command has exited 0
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-extras --streams 1 > gpurun_out/ll_plain.json 2> gpurun_out/ll_plain.err || { echo plain failed; tail -3 gpurun_out/ll_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2500 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-extras --streams 1 > gpurun_out/ll_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches.csv
python scripts/summarise_launches.py gpurun_out/launches.csv gpurun_out/launch_summary.json
