"""Generates tests/golden/tta_*.npz and tests/golden/eval_match.npz by running the REAL reference on CPU
(SURVEY.md 8(f) rows 3-4).  Build container only (the GPU box has no /root/reference):

    python -m oracle.make_golden_tta

* tta_tiny: Model.forward(x, augment=True) (models/yolo.py:582-597) of a fused yolov7-tiny + per-image NMS.
* tta_ensemble: Ensemble.forward (models/experimental.py:74-81) of two differently seeded yolov7-tiny models.
* tta_scale_img: scale_img (utils/torch_utils.py:247-257) on flipped / unflipped inputs at both TTA ratios.
* eval_match: test.py's per-image statistics.  test.py:143-144 and 179-210 are not a callable function, so those
  source lines are read from /root/reference/test.py at run time and exec'd on prepared locals (nothing is copied
  into this repository); box_iou / scale_coords / xywh2xyxy are the reference's own.
Inputs are regenerated from seeds on the test side (oracle/synth.py, ``eval_case`` below).
"""
import os
import sys
import textwrap

import numpy as np
import torch

from .make_golden import GOLD, ROOT, dets_to_np, import_reference


def eval_case(seed, B=4, max_det=60, nc=6, img1=(384, 640)):
    """Seeded detections + labels for the matching fixture: (dets (B, max_det, 6), counts, targets (T, 6), metas).
    Predictions are jittered copies of the labels (so IoUs straddle the 0.5..0.95 thresholds, several predictions
    compete for one target, exact duplicates create IoU ties) plus false positives; image 2 has no labels, image 3 no
    predictions."""
    g = torch.Generator().manual_seed(seed)
    dets = torch.zeros((B, max_det, 6))
    counts = torch.zeros(B, dtype=torch.int32)
    targets, metas = [], []
    for b in range(B):
        h0, w0 = int(torch.randint(300, 900, (1,), generator=g)), int(torch.randint(300, 900, (1,), generator=g))
        gain = min(img1[0] / h0, img1[1] / w0)
        pad = ((img1[1] - w0 * gain) / 2, (img1[0] - h0 * gain) / 2)
        metas.append(dict(shape0=(h0, w0), ratio=(gain, gain), pad=pad))
        nl = 0 if b == 2 else int(torch.randint(5, 25, (1,), generator=g))
        cx = torch.rand(nl, generator=g) * (img1[1] - 2 * pad[0]) + pad[0]
        cy = torch.rand(nl, generator=g) * (img1[0] - 2 * pad[1]) + pad[1]
        wh = torch.rand((nl, 2), generator=g) * 80 + 20
        cls = torch.randint(0, nc, (nl,), generator=g).float()
        if nl > 4:                      # two identical labels: IoU ties -> first-maximum rule
            cx[1], cy[1], wh[1], cls[1] = cx[0], cy[0], wh[0], cls[0]
        lab = torch.stack((torch.full((nl,), float(b)), cls, cx, cy, wh[:, 0], wh[:, 1]), 1)
        targets.append(lab)
        if b == 3:
            continue
        rows = []
        for rep in range(3):            # three noisy predictions per label, growing noise
            noise = (torch.rand((nl, 4), generator=g) - 0.5) * (4.0 + 10.0 * rep)
            x1, y1 = cx - wh[:, 0] / 2 + noise[:, 0], cy - wh[:, 1] / 2 + noise[:, 1]
            x2, y2 = cx + wh[:, 0] / 2 + noise[:, 2], cy + wh[:, 1] / 2 + noise[:, 3]
            pc = torch.where(torch.rand(nl, generator=g) < 0.85, cls, torch.randint(0, nc, (nl,), generator=g).float())
            rows.append(torch.stack((x1, y1, x2, y2, torch.rand(nl, generator=g), pc), 1))
        nf = 8
        fp = torch.rand((nf, 2), generator=g) * torch.tensor([img1[1] - 60.0, img1[0] - 60.0])
        rows.append(torch.cat((fp, fp + torch.rand((nf, 2), generator=g) * 50 + 10, torch.rand((nf, 1), generator=g),
                               torch.randint(0, nc, (nf, 1), generator=g).float()), 1))
        r = torch.cat(rows, 0)
        r = r[r[:, 4].argsort(descending=True)][:max_det]      # NMS output order
        dets[b, :r.shape[0]] = r
        counts[b] = r.shape[0]
    return dets, counts, torch.cat(targets, 0), metas


def _reference_match_lines():
    with open('/root/reference/test.py') as f:
        src = f.read().splitlines()
    predn = textwrap.dedent('\n'.join(src[142:144]))       # test.py:143-144
    match = textwrap.dedent('\n'.join(src[178:208]))       # test.py:179-208
    return compile(predn, 'test.py:143-144', 'exec'), compile(match, 'test.py:179-208', 'exec')


def main():
    sys.path.insert(0, ROOT)
    from oracle.synth import seeded_state_dict, synth_image
    from yolo_series_b200 import zoo
    RefModel, ref_nms = import_reference()
    from models.experimental import Ensemble
    from utils.general import box_iou, scale_coords, xywh2xyxy
    from utils.torch_utils import scale_img
    os.makedirs(GOLD, exist_ok=True)
    with torch.no_grad():
        cfg = zoo.get_cfg('yolov7-tiny')
        models = []
        for seed in (0, 5):
            ref = RefModel(cfg).eval()
            ref.load_state_dict(seeded_state_dict(ref, seed=seed))
            ref.fuse()
            models.append(ref)
        x = synth_image(2, 128, 160, seed=2)
        pred = models[0](x, augment=True)[0]
        dets = [ref_nms(pred[i:i + 1], 0.25, 0.45)[0] for i in range(pred.shape[0])]
        counts, flat = dets_to_np(dets)
        np.savez_compressed(os.path.join(GOLD, 'tta_tiny.npz'), pred=pred.numpy(), counts=counts, dets=flat)
        print('tta_tiny', tuple(pred.shape), counts.tolist())
        ens = Ensemble()
        ens.extend(models)
        pe = ens(x)[0]
        dets = [ref_nms(pe[i:i + 1], 0.25, 0.45)[0] for i in range(pe.shape[0])]
        counts, flat = dets_to_np(dets)
        np.savez_compressed(os.path.join(GOLD, 'tta_ensemble.npz'), pred=pe.numpy(), counts=counts, dets=flat)
        print('tta_ensemble', tuple(pe.shape), counts.tolist())
        xs = synth_image(2, 96, 136, seed=7)
        out = {}
        for ratio in (0.83, 0.67):
            for fi in (None, 2, 3):
                out[f'r{ratio}_f{fi}'] = scale_img(xs.flip(fi) if fi else xs, ratio, gs=32).numpy()
        out['r0.5_same'] = scale_img(xs, 0.5, same_shape=True).numpy()
        np.savez_compressed(os.path.join(GOLD, 'tta_scale_img.npz'), **out)
        print('tta_scale_img', {k: v.shape for k, v in out.items()})

        # ---- evaluation matching: the reference's own lines on seeded detections / labels
        predn_code, match_code = _reference_match_lines()
        img1 = (384, 640)
        d, c, targets, metas = eval_case(31, img1=img1)
        iouv = torch.linspace(0.5, 0.95, 10)
        B, max_det = d.shape[:2]
        correct_all = np.zeros((B, max_det, 10), np.uint8)
        img = torch.zeros((B, 3) + img1)
        shapes = [(m['shape0'], (m['ratio'], m['pad'])) for m in metas]
        for si in range(B):
            pr = d[si, :int(c[si])]
            labels = targets[targets[:, 0] == si, 1:]
            nl = len(labels)
            if len(pr) == 0:
                continue
            ns = dict(pred=pr, labels=labels, nl=nl, niou=10, iouv=iouv, device='cpu', img=img, si=si, shapes=shapes,
                      plots=False, torch=torch, box_iou=box_iou, scale_coords=scale_coords, xywh2xyxy=xywh2xyxy)
            exec(predn_code, ns)
            exec(match_code, ns)
            correct_all[si, :len(pr)] = ns['correct'].numpy()
        k = box_iou(d[0, :int(c[0]), :4], xywh2xyxy(targets[targets[:, 0] == 0, 2:6]))
        np.savez_compressed(os.path.join(GOLD, 'eval_match.npz'), correct=correct_all, box_iou=k.numpy())
        print('eval_match correct per image', correct_all.sum((1, 2)).tolist(), 'iou', tuple(k.shape))


if __name__ == '__main__':
    main()
