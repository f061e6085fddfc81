"""Generates tests/golden/*.npz by running the REAL reference (imported from /root/reference) on CPU.
Run in the build container only (the GPU box has no /root/reference):

    python -m oracle.make_golden

The fixtures pin the oracle (oracle/ref_forward.py, oracle/nms_ref.py) and, through it, the CUDA
path.  Import recipe: SURVEY.md A.4 (matplotlib / seaborn stubs; the mount is read-only).
Inputs are never stored: they are regenerated from seeds by oracle/synth.py on both sides.
"""
import logging
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, 'tests', 'golden')

# (name, B, H, W, conf, iou) -- forward fixtures
FWD_CASES = [
    ('yolov7-tiny', 1, 128, 160, 0.25, 0.45),
    ('yolov7', 1, 64, 64, 0.001, 0.65),
    ('yolov7-w6-custom', 1, 128, 128, 0.001, 0.65),
    ('yolov7-e6e', 1, 128, 128, 0.001, 0.65),
]
# BASELINE.json configs[0]: yolov7-tiny 1x3x640x640, detect.py defaults
E2E_CASE = ('yolov7-tiny', 1, 640, 640, 0.25, 0.45)
# BASELINE.json configs[1]-[3] at their REAL resolutions (round 2): the reference's own models.yolo.Model forward
# (yolo.py:581-631) + non_max_suppression; stored as detections + every 37th row of pred (the npz stays small)
FULL_CASES = [
    ('yolov7', 2, 640, 640, 0.001, 0.65),
    ('yolov7-w6-custom', 1, 1280, 1280, 0.001, 0.65),
    ('yolov7-e6e', 1, 1280, 1280, 0.001, 0.65),
]

# NMS fixtures: (tag, B, A, nc, seed, p_obj, kwargs)
NMS_CASES = [
    ('best', 2, 3000, 80, 11, 0.3, dict(conf_thres=0.001, iou_thres=0.65)),
    ('multi', 2, 1500, 80, 12, 0.3, dict(conf_thres=0.05, iou_thres=0.65, multi_label=True)),
    ('classes', 2, 3000, 80, 13, 0.3, dict(conf_thres=0.25, iou_thres=0.45, classes=[0, 17, 63])),
    ('agnostic', 2, 3000, 80, 14, 0.3, dict(conf_thres=0.25, iou_thres=0.45, agnostic=True)),
    ('nc1', 2, 2000, 1, 15, 0.5, dict(conf_thres=0.25, iou_thres=0.45)),
    ('empty', 2, 500, 80, 16, 0.0, dict(conf_thres=0.25, iou_thres=0.45)),
    ('cut', 1, 40000, 4, 17, 1.0, dict(conf_thres=0.001, iou_thres=0.65)),   # > max_nms candidates
]

# Cross-class suppression (general.py:702-704: the 4096-px class offset does not separate boxes larger than 4096 px):
# (tag, B, A, nc, seed, mixed, kwargs) over oracle.synth.synth_pred_wide
WIDE_CASES = [
    ('wide', 2, 3000, 80, 31, False, dict(conf_thres=0.1, iou_thres=0.05)),
    ('wide_mixed', 2, 3000, 80, 32, True, dict(conf_thres=0.1, iou_thres=0.05)),
]


def import_reference():
    os.environ['PYTHONDONTWRITEBYTECODE'] = '1'
    sys.dont_write_bytecode = True
    for n in ('matplotlib', 'matplotlib.pyplot', 'seaborn'):
        sys.modules.setdefault(n, types.ModuleType(n))
    sys.modules['matplotlib'].rc = lambda *a, **k: None
    sys.modules['matplotlib'].use = lambda *a, **k: None
    sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    if '/root/reference' not in sys.path:
        sys.path.insert(0, '/root/reference')
    logging.disable(logging.CRITICAL)
    from models.yolo import Model
    from utils.general import non_max_suppression
    return Model, non_max_suppression


def dets_to_np(dets):
    counts = np.array([d.shape[0] for d in dets], dtype=np.int64)
    flat = torch.cat([d.reshape(-1, 6) for d in dets], 0).numpy().astype(np.float32) if len(dets) else np.zeros((0, 6), np.float32)
    return counts, flat


def main():
    sys.path.insert(0, ROOT)
    from oracle.synth import seeded_state_dict, synth_image, synth_pred, synth_pred_wide
    from yolo_series_b200 import zoo
    RefModel, ref_nms = import_reference()
    os.makedirs(GOLD, exist_ok=True)
    with torch.no_grad():
        only_full = '--full-only' in sys.argv
        for name, B, H, W, conf, iou in ([] if only_full else FWD_CASES + [E2E_CASE]) + ([] if '--wide' in sys.argv else FULL_CASES):
            cfg = zoo.get_cfg(name)
            ref = RefModel(cfg).eval()
            sd = seeded_state_dict(ref, seed=0)
            ref.load_state_dict(sd)
            ref.fuse()
            x = synth_image(B, H, W, seed=2)
            pred, raws = ref(x)
            dets = [ref_nms(pred[i:i + 1], conf, iou)[0] for i in range(B)]   # per image: no time_limit drop
            counts, flat = dets_to_np(dets)
            tag = f'{name}_{H}x{W}'
            full = max(H, W) < 640
            out = dict(counts=counts, dets=flat, conf=conf, iou=iou, shape=np.array(pred.shape),
                       cand=np.array([(pred[..., 4] > conf).sum().item()]))
            if full:
                out['pred'] = pred.numpy()
                out['raw0'] = raws[0].numpy()
            else:
                out['pred_sub'] = pred[:, ::37].numpy()
            np.savez_compressed(os.path.join(GOLD, f'fwd_{tag}.npz'), **out)
            print(tag, 'pred', tuple(pred.shape), 'candidates', int(out['cand'][0]), 'dets', counts.tolist())
        for tag, B, A, nc, seed, p_obj, kw in ([] if only_full else NMS_CASES):
            pred = synth_pred(B, A, nc, seed=seed, p_obj=p_obj)
            dets = [ref_nms(pred[i:i + 1].clone(), **kw)[0] for i in range(B)]
            counts, flat = dets_to_np(dets)
            np.savez_compressed(os.path.join(GOLD, f'nms_{tag}.npz'), counts=counts, dets=flat)
            print('nms', tag, counts.tolist())
        for tag, B, A, nc, seed, mixed, kw in ([] if (only_full and '--wide' not in sys.argv) else WIDE_CASES):
            pred = synth_pred_wide(B, A, nc, seed=seed, mixed=mixed)
            dets = [ref_nms(pred[i:i + 1].clone(), **kw)[0] for i in range(B)]
            counts, flat = dets_to_np(dets)
            np.savez_compressed(os.path.join(GOLD, f'nms_{tag}.npz'), counts=counts, dets=flat)
            print('nms', tag, counts.tolist())


if __name__ == '__main__':
    main()
