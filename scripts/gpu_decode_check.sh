#!/bin/bash
timeout 900 python -m pytest -q -m gpu --timeout 300 -p no:cacheprovider tests/test_gpu_nms.py tests/test_gpu_e2e.py tests/test_gpu_tta.py tests/test_gpu_fullsize.py -x 2>&1 | tail -5
python scripts/tta_breakdown.py 2>&1 | grep "forward(augment)\|detect"
python - <<'PY'
import torch, sys
sys.path.insert(0,'.')
from yolo_series_b200 import Model
from yolo_series_b200.synth import seeded_state_dict, synth_image
m = Model('yolov7'); m.load_state_dict(seeded_state_dict(m,0)); m = m.cuda().eval().half()
x = synth_image(32,640,640,seed=1,dtype=torch.uint8).cuda()
def t(fn,n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True); a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b)/n
tt = t(lambda: m.run_trunk(x)); tf = t(lambda: m(x))
print(f'trunk {tt:.3f} ms, forward (trunk + decode with raw) {tf:.3f} ms, decode {tf-tt:.3f} ms; bytes 32*25200*85*(4+4+4) = {32*25200*85*12/1e6:.0f} MB -> {32*25200*85*12/1e6/(tf-tt):.0f} GB/s')
PY
