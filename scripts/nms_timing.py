"""Phase timestamps of k_sort_nms (globaltimer): build nms.cu with -DYB_NMS_TIMING first, e.g.
  cd yolo_series_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC -DYB_NMS_TIMING -c nms.cu -o build/nms.o && make
The kernel then writes 64 timestamps per image into the keys_b workspace (unused unless a band overflows)."""
import sys, torch
sys.path.insert(0, '.')
from yolo_series_b200 import Model, nms
from yolo_series_b200.synth import seeded_state_dict, synth_image
import ctypes as C
from yolo_series_b200 import _lib
for name, B in (('yolov7', 32), ('yolov7-tiny', 1)):
    m = Model(name); m.load_state_dict(seeded_state_dict(m, 0)); m = m.cuda().eval().half()
    x = synth_image(B, 640, 640, seed=100, dtype=torch.uint8).cuda()
    for _ in range(3):
        out, meta = m.detect(x, 0.001, 0.65, packed=True)
    torch.cuda.synchronize()
    ws = list(nms._ws_cache.values())[0]
    A, no = 25200, 85
    # locate keys_b: carve = count(B*4) | keys_a (B*cap*8) | keys_b ...
    al = lambda v: (v + 255) // 256 * 256
    base = (ws.data_ptr() + 255) // 256 * 256 - ws.data_ptr()
    off = base + al(B * 4) + al(B * A * 8)
    t = ws[off: off + B * 64 * 8].view(torch.int64).view(B, 64).cpu()
    t0 = t[:, 0:1]
    rel = (t - t0).float() / 1000.0
    print(name, 'B', B, 'n cand/img (count):', ws[base: base + 4 * B].view(torch.int32)[:4].tolist())
    r = rel[0]
    print(' select_band %.1f us | compaction %.1f | band sort %.1f' % (r[1], r[2] - r[1], r[3] - r[2]))
    ch = [float(r[4 + i]) for i in range(3) if t[0, 4 + i] > t[0, 0]]
    print(' chunk starts (us):', [round(c, 1) for c in ch], 'end', round(float(r[40]), 1))
    print(' chunk0: load %.1f | A %.1f | B %.1f | C %.1f | D.. %.1f' % (r[50]-r[4], r[51]-r[50], r[52]-r[51], r[53]-r[52], r[5]-r[53]))
    nms._ws_cache.clear()
