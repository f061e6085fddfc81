#!/bin/bash
# One compute-sanitizer tool per gpurun call (B200_PROFILING.md), on small cases, only after the same command has
# passed without the tool.  TOOL=memcheck|racecheck|synccheck|initcheck  SEL="<pytest args>"
TOOL=${TOOL:-memcheck}
SEL=${SEL:-"tests/test_gpu_nms.py"}
mkdir -p gpurun_out
PT="python -m pytest -q -m gpu -p no:cacheprovider -x $SEL"
timeout 600 $PT > gpurun_out/san_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/san_plain.log; exit 1; }
tail -1 gpurun_out/san_plain.log
timeout 2400 compute-sanitizer --tool $TOOL --error-exitcode 7 --print-limit 20 $PT > gpurun_out/san_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|hazard" gpurun_out/san_$TOOL.log | sort | uniq -c | head -20
