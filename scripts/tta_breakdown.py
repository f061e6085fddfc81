"""Where the time of Model.detect(augment=True) goes (yolov7 640 B=32 bf16): CUDA events per stage."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yolo_series_b200 import Model, _lib                                  # noqa: E402
from yolo_series_b200.nms import _run_nms                                 # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image        # noqa: E402

m = Model('yolov7')
m.load_state_dict(seeded_state_dict(m, 0))
m = m.cuda().eval().half()
B = 32
x = synth_image(B, 640, 640, seed=1, dtype=torch.uint8).cuda()


def t(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


passes = m._passes(640, 640, True)
print('passes', passes)
xs = []
for si, fi, hw, HW in passes:
    xi = x if si == 1 else m._scaled_input(x, si, fi, hw, HW)
    xs.append(xi)
    if si != 1:
        print(f'scale_img {si}: {t(lambda: m._scaled_input(x, si, fi, hw, HW)):.3f} ms')
    print(f'trunk {HW} ({xi.dtype}): {t(lambda: m.run_trunk(xi)):.3f} ms')
    xu = synth_image(B, HW[0], HW[1], seed=1, dtype=torch.uint8).cuda()
    print(f'trunk {HW} (uint8 input): {t(lambda: m.run_trunk(xu)):.3f} ms')
pred = m(x, augment=True)[0]
print('pred', tuple(pred.shape))
print(f'forward(augment) total: {t(lambda: m(x, augment=True)):.3f} ms')
print(f'nms on TTA pred: {t(lambda: _run_nms(pred, B, pred.shape[1], 85, 0.001, 0.65, None, False, False, 30000, 300, pred.device, return_packed=True)):.3f} ms')
print(f'detect(augment) total: {t(lambda: m.detect(x, 0.001, 0.65, augment=True, packed=True)):.3f} ms')
print(f'detect plain: {t(lambda: m.detect(x, 0.001, 0.65, packed=True)):.3f} ms')
