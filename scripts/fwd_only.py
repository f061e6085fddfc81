import sys, torch
sys.path.insert(0, '.')
from yolo_series_b200 import Model
from yolo_series_b200.synth import seeded_state_dict, synth_image
m = Model('yolov7'); m.load_state_dict(seeded_state_dict(m, 0)); m = m.cuda().eval().half()
x = synth_image(32, 640, 640, seed=1, dtype=torch.uint8).cuda()
for _ in range(3):
    p = m(x)
torch.cuda.synchronize()
print('ok', tuple(p[0].shape))
