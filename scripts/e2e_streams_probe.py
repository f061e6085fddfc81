"""detect_batches with one vs two compute streams, alternating, long runs (same process, same box)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import Model  # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
STEPS = int(os.environ.get('STEPS', 200))
m = Model('yolov7')
m.load_state_dict(seeded_state_dict(m, 0))
m = m.cuda().eval().half()
xh = synth_image(B, 640, 640, seed=1, dtype=torch.uint8).pin_memory()
xd = xh.cuda()


def run(n, cs):
    last = None
    for last in m.detect_batches((xh for _ in range(n)), 0.001, 0.65, max_det=300, compute_streams=cs):
        pass
    return last


def run_dev(n, cs):
    side = m.compute_streams(2)
    for k in range(n):
        if cs == 1:
            m.detect(xd, 0.001, 0.65, packed=True)
        else:
            with torch.cuda.stream(side[k & 1]):
                m.detect(xd, 0.001, 0.65, packed=True)
    torch.cuda.synchronize()


for cs in (1, 2):
    run(4, cs), run_dev(4, cs)
torch.cuda.synchronize()
for rep in range(3):
    for fn, tag in ((run, 'e2e'), (run_dev, 'dev')):
        for cs in (1, 2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn(STEPS, cs)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            print(f'{tag} compute_streams={cs}: {STEPS * B / dt:.1f} img/s  {dt / STEPS * 1e3:.4f} ms/step', flush=True)
