"""Where the time of Model.forward(augment=True) + NMS goes (yolo.py:582-597): per pass scale_img, trunk, decode; then NMS."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import Model, non_max_suppression  # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
m = Model('yolov7')
m.load_state_dict(seeded_state_dict(m, 0))
m = m.cuda().eval().half()
x = synth_image(B, 640, 640, seed=1, dtype=torch.uint8).cuda()


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


t_plain, _ = timed(lambda: m.detect(x, 0.001, 0.65, packed=True))
t_fwd, pa = timed(lambda: m(x, augment=True)[0])
t_all, _ = timed(lambda: m.detect(x, 0.001, 0.65, packed=True, augment=True))
print(f'plain step {t_plain:.3f} ms; TTA forward (3 passes -> pred {tuple(pa.shape)}) {t_fwd:.3f} ms; TTA forward + NMS {t_all:.3f} ms')
H = W = 640
for si, fi, hw, HW in m._passes(H, W, True):
    xi = x if (si == 1.0 and not fi) else m._scaled_input(x, si, fi, hw, HW)
    t_s, _ = timed(lambda: None if (si == 1.0 and not fi) else m._scaled_input(x, si, fi, hw, HW))
    t_t, eng = timed(lambda: m.run_trunk(xi))
    print(f'  pass scale {si} flip {fi} -> {HW}: scale_img {t_s:.3f} ms, trunk {t_t:.3f} ms ({eng.conv_flops / t_t / 1e9:.0f} TFLOP/s)')
t_n, _ = timed(lambda: non_max_suppression(pa, 0.001, 0.65))
print(f'  stand-alone NMS on {pa.shape[1]} rows/image (list form, one host sync): {t_n:.3f} ms')
