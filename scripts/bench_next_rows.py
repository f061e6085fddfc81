"""Measures the SURVEY.md 8(f) "next" rows on one B200: letterbox / rescale (rows 1-2), TTA + ensemble (row 3),
box_iou / match (row 4).  Device times by CUDA events after warm-up; algorithmic bytes stated per kernel; the oracle
restatement of each row is timed beside it on the host cores (bounded sample).  Prints one JSON object."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.eval_ref import box_iou_ref, match_image_ref           # noqa: E402
from oracle.make_golden_tta import eval_case                        # noqa: E402
from oracle.synth import seeded_state_dict                          # noqa: E402
from oracle.tta_ref import scale_img_ref                            # noqa: E402
from yolo_series_b200 import Ensemble, Model                        # noqa: E402
from yolo_series_b200.evaluate import box_iou, match_batch          # noqa: E402
from yolo_series_b200.preprocess import letterbox_batch, rescale_detections  # noqa: E402

PEAK = 6545.6
try:
    PEAK = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs'])
except Exception:
    pass


def dev_ms(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def host_s(fn):
    t = time.perf_counter()
    fn()
    return time.perf_counter() - t


def main():
    torch.cuda.set_device(0)
    out = {'hbm_peak_GBps': PEAK, 'cores': os.cpu_count()}
    g = torch.Generator().manual_seed(0)
    B = 32
    # ---- row 1: letterbox + BGR->RGB + HWC->CHW, 32 frames 1080x1920 -> 640x640 (384x640 content + borders)
    frames = [torch.randint(0, 256, (1080, 1920, 3), generator=g, dtype=torch.uint8).cuda() for _ in range(B)]
    dst = torch.empty((B, 3, 640, 640), dtype=torch.uint8, device='cuda')
    ms = dev_ms(lambda: letterbox_batch(frames, 640, auto=False, out=dst))
    # linear resize by 1/3 touches a 2x2 footprint out of every 3x3 source block: algorithmic = rows actually read
    by = B * (1080 * 1920 * 3 * (2 / 3) + 3 * 640 * 640)
    out['letterbox'] = {'ms': ms, 'frames': B, 'us_per_frame': ms * 1e3 / B, 'algorithmic_GBps': by / ms / 1e6,
                        'frac_of_hbm_peak': by / ms / 1e6 / PEAK,
                        'bytes': 'B*(2/3 of the 1080x1920x3 source rows + 3x640x640 out)'}
    import cv2
    f0 = frames[0].cpu().numpy()
    t = host_s(lambda: [cv2.copyMakeBorder(cv2.resize(f0, (640, 360), interpolation=cv2.INTER_LINEAR), 140, 140, 0, 0,
                                           cv2.BORDER_CONSTANT, value=(114, 114, 114))[:, :, ::-1].transpose(2, 0, 1).copy()
                        for _ in range(32)])
    out['letterbox']['cpu_us_per_frame'] = t / 32 * 1e6
    # ---- row 2: scale_coords + clip + round on the packed NMS output
    dets = torch.rand((B, 300, 6), generator=g).cuda() * 640
    cnt = torch.full((B,), 300, dtype=torch.int32, device='cuda')
    metas = [dict(shape0=(1080, 1920), ratio=(1 / 3, 1 / 3), pad=(0.0, 140.0)) for _ in range(B)]
    ms = dev_ms(lambda: rescale_detections(dets, cnt, (640, 640), metas))
    out['rescale_detections'] = {'ms': ms, 'rows': B * 300, 'note': 'launch-latency bound (230 KB)'}
    # ---- row 3: scale_img
    x = torch.rand((B, 3, 640, 640), generator=g).cuda()
    m = Model('yolov7')
    m.load_state_dict(seeded_state_dict(m, seed=0))
    m = m.cuda().eval()
    ms = dev_ms(lambda: m._scaled_input(x, 0.83, 3, (531, 531), (544, 544)))
    by = B * 3 * 4 * (640 * 640 + 544 * 544)
    out['scale_img'] = {'ms': ms, 'algorithmic_GBps': by / ms / 1e6, 'frac_of_hbm_peak': by / ms / 1e6 / PEAK,
                        'bytes': 'B*3*4*(640*640 in + 544*544 out)', 'shape': [B, 3, 640, 640], 'ratio': 0.83}
    xc = x[:4].cpu()
    t = host_s(lambda: scale_img_ref(xc, 0.83, gs=32, flip=3))
    out['scale_img']['cpu_port_ms_per_image'] = t / 4 * 1e3
    t = host_s(lambda: torch.nn.functional.interpolate(xc.flip(3), size=(531, 531), mode='bilinear', align_corners=False))
    out['scale_img']['cpu_torch_ms_per_image'] = t / 4 * 1e3
    # ---- row 3: TTA end to end, yolov7 640 B=32 bf16, conf 0.001 / iou 0.65 (3 passes: 640, 544 flipped, 448)
    xu = torch.randint(0, 256, (B, 3, 640, 640), generator=g, dtype=torch.uint8).cuda()
    ms1 = dev_ms(lambda: m.detect(xu, 0.001, 0.65, packed=True), iters=10)
    msa = dev_ms(lambda: m.detect(xu, 0.001, 0.65, packed=True, augment=True), iters=10)
    out['tta'] = {'model': 'yolov7 640 B=32 bf16', 'plain_ms': ms1, 'augment_ms': msa, 'augment_images_per_s': B / msa * 1e3,
                  'rows_per_image': m.rows_per_image(640, 640, True), 'passes': '640 | 544 lr-flipped | 448',
                  'flops_ratio_expected': 1 + (544 / 640) ** 2 + (448 / 640) ** 2, 'time_ratio': msa / ms1}
    # ---- row 3: ensemble of two yolov7-tiny
    tiny = []
    for s in (0, 5):
        t_ = Model('yolov7-tiny')
        t_.load_state_dict(seeded_state_dict(t_, seed=s))
        tiny.append(t_.cuda().eval())
    ens = Ensemble(tiny)
    mse = dev_ms(lambda: ens(xu), iters=10)
    ms0 = dev_ms(lambda: tiny[0](xu), iters=10)
    out['ensemble'] = {'members': '2 x yolov7-tiny 640 B=32 bf16', 'ms': mse, 'single_member_ms': ms0}
    # ---- row 4: box_iou
    a = torch.rand((4096, 4), generator=g) * 600
    a[:, 2:] += a[:, :2]
    b = torch.rand((4096, 4), generator=g) * 600
    b[:, 2:] += b[:, :2]
    ac, bc = a.cuda(), b.cuda()
    ms = dev_ms(lambda: box_iou(ac, bc))
    by = 4096 * 4096 * 4
    out['box_iou'] = {'ms': ms, 'n': 4096, 'm': 4096, 'algorithmic_GBps': by / ms / 1e6,
                      'frac_of_hbm_peak': by / ms / 1e6 / PEAK, 'bytes': 'n*m*4 written once'}
    t = host_s(lambda: box_iou_ref(a.numpy(), b.numpy()))
    out['box_iou']['cpu_port_ms'] = t * 1e3
    # ---- row 4: matching, B=32 x 300 detections
    d, c, targets, metas = eval_case(2, B=32, max_det=300, nc=80, img1=(384, 640))
    dc, cc, tc = d.cuda(), c.cuda(), targets.cuda()
    iouv = torch.linspace(0.5, 0.95, 10).cuda()
    ms = dev_ms(lambda: match_batch(dc, cc, tc, (384, 640), metas, iouv))
    out['match_batch'] = {'ms': ms, 'images': 32, 'max_det': 300, 'targets': int(targets.shape[0]),
                          'us_per_image': ms * 1e3 / 32, 'note': 'host shim (table upload) + 2 launches; latency bound'}

    def cpu_match():
        for i in range(8):
            lab = targets[targets[:, 0] == i, 1:].numpy()
            mm = metas[i]
            match_image_ref(d[i, :int(c[i])].numpy(), lab, mm['ratio'][0], mm['pad'], mm['shape0'], iouv.cpu().numpy())
    t = host_s(cpu_match)
    out['match_batch']['cpu_port_us_per_image'] = t / 8 * 1e6
    print(json.dumps(out))


if __name__ == '__main__':
    main()
