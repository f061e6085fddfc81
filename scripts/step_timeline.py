"""Timeline of ONE graph replay of the trunk from per-CTA globaltimer stamps of every tcgen05 conv launch (env
YB_CONV_TS + YB_CONV_TS_STRIDE, read at plan creation: each plan gets its own slot in creation = op order).  Per launch:
first CTA entry, griddepcontrol.wait passed, last CTA exit -- and the gap to the next stamped launch.
Usage: python scripts/step_timeline.py [model] [B] [HW] > gpurun_out/step_timeline.tsv"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'yolov7'
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
HW = int(sys.argv[3]) if len(sys.argv) > 3 else 640
CTAS, SLOTS = 320, 16
stride = CTAS * SLOTS
ts = torch.zeros(400 * stride, dtype=torch.int64, device='cuda')
os.environ['YB_CONV_TS'] = hex(ts.data_ptr())
os.environ['YB_CONV_TS_STRIDE'] = str(stride)

from yolo_series_b200 import Model  # noqa: E402
from yolo_series_b200.synth import seeded_state_dict, synth_image  # noqa: E402

m = Model(name)
m.load_state_dict(seeded_state_dict(m, 0))
m = m.cuda().eval().half()
x = synth_image(B, HW, HW, seed=1, dtype=torch.uint8).cuda()
eng = m.run_trunk(x)
for _ in range(3):
    m.run_trunk(x)
torch.cuda.synchronize()
ts.zero_()
torch.cuda.synchronize()
m.run_trunk(x)                      # the stamped replay
torch.cuda.synchronize()
t = ts.view(-1, CTAS, SLOTS).cpu()
stamped = [i for i in range(len(eng.ops)) if eng.op_kernel(i) in ('k_conv_tc', 'k_conv3_tc', 'k_conv3s2p_tc')]
rows = []
for slot, i in enumerate(stamped):
    a = t[slot]
    a = a[a[:, 0] > 0]
    if a.shape[0] == 0:
        continue
    wait = a[:, 2][a[:, 2] > 0]
    ex = a[:, 9][a[:, 9] > 0]
    rows.append((i, eng.ops[i].name, eng.op_kernel(i), a.shape[0], int(a[:, 0].min()), int(wait.min()) if wait.numel() else 0,
                 int(ex.median()) if ex.numel() else 0, int(ex.max()) if ex.numel() else 0))
t0 = rows[0][4]
print('op\tname\tkernel\tctas\tentry_us\twait_done_us\texit_median_us\texit_max_us\tspan_us(wait->exit)\ttail_us(median->max exit)\tgap_to_next_wait_us')
tot_span = tot_gap = tot_tail = 0.0
for k, r in enumerate(rows):
    i, nm, kern, n, e, w, xm, xx = r
    nxt = rows[k + 1] if k + 1 < len(rows) else None
    adjacent = nxt is not None and nxt[0] == i + 1          # the next op is a stamped conv too
    gap = (nxt[5] - xx) / 1e3 if adjacent else float('nan')
    span = (xx - w) / 1e3
    tot_span += span
    tot_tail += (xx - xm) / 1e3
    if adjacent:
        tot_gap += gap
    print(f'{i}\t{nm}\t{kern}\t{n}\t{(e - t0) / 1e3:.2f}\t{(w - t0) / 1e3:.2f}\t{(xm - t0) / 1e3:.2f}\t{(xx - t0) / 1e3:.2f}\t{span:.2f}\t{(xx - xm) / 1e3:.2f}\t{gap:.2f}')
print(f'# {len(rows)} stamped launches: sum of spans {tot_span:.1f} us, of exit tails {tot_tail:.1f} us, of conv->conv gaps {tot_gap:.1f} us; '
      f'first entry to last exit {(rows[-1][7] - t0) / 1e3:.1f} us')
