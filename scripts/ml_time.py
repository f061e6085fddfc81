"""Timing of the multi_label=True variant of BASELINE configs[1] (test.py:126 calls NMS with multi_label=True)."""
import sys, torch
sys.path.insert(0, '.')
from yolo_series_b200 import Model
from yolo_series_b200.synth import seeded_state_dict, synth_image
m = Model('yolov7'); m.load_state_dict(seeded_state_dict(m, 0)); m = m.cuda().eval().half()
x = synth_image(32, 640, 640, seed=100, dtype=torch.uint8).cuda()
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True); a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / n
tt = t(lambda: m.run_trunk(x))
for ml in (False, True):
    ms = t(lambda: m.detect(x, 0.001, 0.65, multi_label=ml, packed=True))
    out, meta = m.detect(x, 0.001, 0.65, multi_label=ml, packed=True)
    print(f'multi_label={ml}: detect {ms:.3f} ms (trunk {tt:.3f}, decode+NMS {ms - tt:.3f}) -> {32 / ms * 1e3:.0f} img/s; kept/img {float(meta[:-1].float().mean()):.0f}')
