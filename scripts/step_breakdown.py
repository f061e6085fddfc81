"""Where one bench step goes: trunk graph replay, fused decode+filter, sort+NMS (yolov7 640, B=32, bf16)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import Model, synth  # noqa: E402

B = int(os.environ.get('YB_B', 32))
NAME = sys.argv[1] if len(sys.argv) > 1 else 'yolov7'
SZ = int(sys.argv[2]) if len(sys.argv) > 2 else 640
torch.manual_seed(0)
model = Model(NAME)
model.load_state_dict(synth.seeded_state_dict(model, seed=0))
model = model.cuda().eval().half()
x = synth.synth_image(B, SZ, SZ, seed=1, dtype=torch.uint8).cuda()


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


t_trunk = timed(lambda: model.run_trunk(x))
t_det = timed(lambda: model.detect(x, 0.001, 0.65, max_det=300, packed=True))
out, meta = model.detect(x, 0.001, 0.65, max_det=300, packed=True)
torch.cuda.synchronize()
print(NAME, SZ, 'B', B, f'{B / t_det * 1e3:.0f} img/s'); print(f'trunk(graph) {t_trunk:.3f} ms   detect {t_det:.3f} ms   decode+nms {t_det - t_trunk:.3f} ms')
print('counts', meta[:B].tolist()[:8], 'overflow', int(meta[B]))
eng = model.run_trunk(x)
print('launches per trunk', eng.num_launches())
prof = eng.profile()
print('sum of per-op events (eager)', sum(p[2] for p in prof))
