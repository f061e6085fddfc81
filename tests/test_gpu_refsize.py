"""BASELINE.json configs[1]-[3] at their REAL resolutions against the REFERENCE ITSELF: the goldens
tests/golden/fwd_{yolov7_640x640, yolov7-w6-custom_1280x1280, yolov7-e6e_1280x1280}.npz were written by
oracle/make_golden.py from the reference's own models.yolo.Model forward (yolo.py:581-631) + non_max_suppression
(general.py:628-720) on the seeded weights / images both sides regenerate.  Stored: the detections and every 37th
row of pred.

Bars (north_star):
  fp32 mode  boxes <= 1e-2 px, scores <= 1e-3 on every stored pred row; detections: every reference row must have an
             IoU >= 0.99 same-class partner, except rows within rounding of a threshold (the max_det cut or a
             suppression decision) -- those are listed and must sit within 2e-3 of the cut score or be explained by a
             partner at IoU >= 0.9 (a near-tie that swapped rank).
  bf16 mode  stated tolerance: p99 box error <= 2 % of the box size + 1 px and p99 score error <= 0.03, AND the maximum
             is bounded (<= 4x those figures); the detection-level agreement with the reference's detections is measured
             and asserted against the floor stated in BASELINE.md.
"""
import json
import os

import numpy as np
import pytest
import torch

from oracle.make_golden import FULL_CASES
from oracle.synth import seeded_state_dict, synth_image

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(name, mode):
    from yolo_series_b200 import Model
    m = Model(name)
    m.load_state_dict(seeded_state_dict(m, seed=0))
    m.set_precision(mode)
    return m.cuda().eval()


def _iou(a, b):
    """(n,4) x (m,4) xyxy -> (n,m), plain fp32 torch on the host (checker arithmetic, not the product's)."""
    lt = torch.max(a[:, None, :2], b[None, :, :2])
    rb = torch.min(a[:, None, 2:], b[None, :, 2:])
    inter = (rb - lt).clamp(min=0).prod(2)
    aa = (a[:, 2] - a[:, 0]) * (a[:, 3] - a[:, 1])
    ab = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    return inter / (aa[:, None] + ab[None, :] - inter).clamp(min=1e-12)


def _matched(ref, got, thr):
    """Greedy one-to-one matching in reference order: same class, IoU >= thr.  Returns a bool mask over ref rows."""
    if ref.shape[0] == 0:
        return torch.zeros(0, dtype=torch.bool)
    if got.shape[0] == 0:
        return torch.zeros(ref.shape[0], dtype=torch.bool)
    iou = _iou(ref[:, :4], got[:, :4])
    iou[ref[:, 5:6] != got[None, :, 5]] = -1
    used = torch.zeros(got.shape[0], dtype=torch.bool)
    ok = torch.zeros(ref.shape[0], dtype=torch.bool)
    for i in range(ref.shape[0]):
        row = iou[i].clone()
        row[used] = -1
        j = int(row.argmax())
        if row[j] >= thr:
            used[j] = True
            ok[i] = True
    return ok


def _split(g):
    counts, flat = g['counts'], torch.from_numpy(g['dets'])
    out, o = [], 0
    for c in counts.tolist():
        out.append(flat[o:o + c])
        o += c
    return out


def _record(tag, d):
    """Measurements go to gpurun_out/ (when it exists) so the numbers quoted in BASELINE.md can be re-read."""
    p = os.path.join(ROOT, 'gpurun_out')
    if os.path.isdir(p):
        with open(os.path.join(p, 'parity_refsize.jsonl'), 'a') as f:
            f.write(json.dumps({'case': tag, **d}) + '\n')
    print(tag, d)


@pytest.mark.parametrize('case', FULL_CASES, ids=[f'{c[0]}_{c[2]}' for c in FULL_CASES])
def test_full_size_fp32_mode_matches_reference(case, gold):
    name, B, H, W, conf, iou = case
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    m = _build(name, 'fp32')
    x = synth_image(B, H, W, seed=2).cuda()
    pred, _ = m(x)
    assert tuple(pred.shape) == tuple(g['shape'])
    sub, ref = pred[:, ::37].cpu(), torch.from_numpy(g['pred_sub'])
    box_err = float((sub[..., :4] - ref[..., :4]).abs().max())
    sc_err = float((sub[..., 4:] - ref[..., 4:]).abs().max())
    assert box_err <= 1e-2 and sc_err <= 1e-3, (box_err, sc_err)
    dets = m.detect(x, conf, iou)
    refd = _split(g)
    worst = 1.0
    for i in range(B):
        got, r = dets[i].cpu(), refd[i]
        assert abs(got.shape[0] - r.shape[0]) <= 2
        strict = _matched(r, got, 0.99)
        rate = float(strict.float().mean())
        worst = min(worst, rate)
        # rows without a strict partner must be explainable by a threshold: the max_det cut (their score is within
        # 2e-3 of the last kept score) or a suppression flip / near-tie (a partner exists at IoU >= 0.9)
        loose = _matched(r, got, 0.9)
        cut = float(r[-1, 4]) if r.shape[0] else 0.0
        for j in (~strict).nonzero().flatten().tolist():
            near_cut = abs(float(r[j, 4]) - cut) <= 2e-3
            assert near_cut or bool(loose[j]), f'image {i} reference row {j} {r[j].tolist()} has no counterpart'
        assert rate >= 0.95, f'image {i}: only {rate:.3f} of the reference detections matched at IoU >= 0.99'
    _record(f'{name}_{H} fp32', {'pred_box_max_err_px': box_err, 'pred_score_max_err': sc_err,
                                 'detections_matched_iou99_min_over_images': worst})


# floor of the IoU>=0.5 same-class agreement of the bf16 detections with the reference's fp32 detections (conf 0.001:
# every image saturates max_det, so rank perturbations at the cut move rows in and out); measured values: BASELINE.md
BF16_AGREE_FLOOR = 0.85


@pytest.mark.parametrize('case', FULL_CASES, ids=[f'{c[0]}_{c[2]}' for c in FULL_CASES])
def test_full_size_bf16_mode_vs_reference(case, gold):
    name, B, H, W, conf, iou = case
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    m = _build(name, 'bf16')
    x = synth_image(B, H, W, seed=2).cuda()
    pred, _ = m(x)
    sub, ref = pred[:, ::37].cpu(), torch.from_numpy(g['pred_sub'])
    size = ref[..., 2:4].clamp(min=1.0).max(-1, keepdim=True)[0]
    rel = ((sub[..., :4] - ref[..., :4]).abs() / (0.02 * size + 1.0)).flatten()
    sc = (sub[..., 4:] - ref[..., 4:]).abs().flatten()
    stats = {'box_err_over_tol_p99': float(rel.quantile(0.99)), 'box_err_over_tol_max': float(rel.max()),
             'score_err_p99': float(sc.quantile(0.99)), 'score_err_max': float(sc.max())}
    dets = m.detect(x, conf, iou)
    refd = _split(g)
    a50, a90 = [], []
    for i in range(B):
        got, r = dets[i].cpu(), refd[i]
        a50.append(float(_matched(r, got, 0.5).float().mean()))
        a90.append(float(_matched(r, got, 0.9).float().mean()))
    stats.update({'detections_matched_iou50': a50, 'detections_matched_iou90': a90})
    _record(f'{name}_{H} bf16', stats)
    assert stats['box_err_over_tol_p99'] <= 1.0 and stats['score_err_p99'] <= 0.03
    assert stats["box_err_over_tol_max"] <= 1.0 and stats["score_err_max"] <= 0.03, stats      # the MAX is inside the stated tolerance too
    assert min(a50) >= BF16_AGREE_FLOOR, stats
