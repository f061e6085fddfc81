#!/bin/bash
# plain run, then one ncu --set full capture of the first-layer kernel (yolov7 stem, B=32, 640x640 u8)
mkdir -p gpurun_out
cat > /tmp/stem_run.py <<'PY'
import sys, os, torch
sys.path.insert(0, os.getcwd())
from yolo_series_b200 import ops
g = torch.Generator().manual_seed(0)
img = torch.randint(0, 256, (32, 3, 640, 640), generator=g, dtype=torch.uint8).cuda()
w = torch.randn((32, 3, 3, 3), generator=g) / 27 ** 0.5
b = torch.randn((32,), generator=g)
s = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for _ in range(3):
    y = ops.stem_conv(img, w, b, s, 1, 0.0, False)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    y = ops.stem_conv(img, w, b, s, 1, 0.0, False)
e1.record(); torch.cuda.synchronize()
print('stem ms', e0.elapsed_time(e1) / 10)
PY
timeout 300 python /tmp/stem_run.py || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stem --launch-skip 3 --launch-count 1 \
  -o gpurun_out/$1 -f python /tmp/stem_run.py > gpurun_out/$1_ncu.log 2>&1
echo "ncu rc=$?"
