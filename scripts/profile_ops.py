"""Per-op device time of one eager trunk pass (CUDA events between ops, yb_engine_profile).
Usage: python scripts/profile_ops.py [model] [B] [HW] > gpurun_out/ops.tsv"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import Model  # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'yolov7'
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
HW = int(sys.argv[3]) if len(sys.argv) > 3 else 640
m = Model(name)
m.load_state_dict(seeded_state_dict(m, 0))
m = m.cuda().eval().half()
x = synth_image(B, HW, HW, seed=1, dtype=torch.uint8).cuda()
eng = m.run_trunk(x)
for _ in range(2):
    eng.profile()
reps = [eng.profile() for _ in range(5)]
L = eng.L
print('op\tkind\tCin\tCout\tk\ts\tHout\tups\tms\tTFLOPs\tGBps_algo')
tot = 0
for i, op in enumerate(eng.ops):
    ms = sorted(r[i][2] for r in reps)[2]
    tot += ms
    fl = reps[0][i][3]
    vin = L.vals[op.ins[0]]
    hin = HW * vin.num // vin.den
    if op.kind == 1:
        fc = op.conv
        ho = hin // fc.s
        es_out = 4 if L.vals[op.outs[0]].f32 else 2
        byts = B * (hin * hin * max(fc.c_in, 16) * 2 + ho * ho * op.upsample ** 2 * fc.c_out * es_out)
        print(f'{op.name}\tconv\t{fc.c_in}\t{fc.c_out}\t{fc.k}\t{fc.s}\t{ho}\t{op.upsample}\t{ms:.4f}\t{fl / ms / 1e9:.1f}\t{byts / ms / 1e6:.0f}')
    else:
        nout = len(op.outs)
        ho = hin // 2 if op.kind == 2 else hin
        byts = B * vin.C * 2 * (hin * hin + nout * ho * ho)
        print(f'{op.name}\tk{op.kind}\t{vin.C}\t{vin.C}\t-\t-\t{ho}\t1\t{ms:.4f}\t-\t{byts / ms / 1e6:.0f}')
print(f'# total {tot:.3f} ms, conv flops {sum(r[3] for r in reps[0]) / 1e12:.3f} TFLOP')
