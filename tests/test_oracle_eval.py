"""Pins the evaluation-matching oracle (oracle/eval_ref.py) against the reference's own lines."""
import numpy as np
import torch

from oracle.eval_ref import box_iou_ref, match_image_ref
from oracle.make_golden_tta import eval_case


def test_match_reproduces_reference(gold):
    g = gold('eval_match.npz')
    dets, counts, targets, metas = eval_case(31, img1=(384, 640))
    iouv = torch.linspace(0.5, 0.95, 10).numpy()
    for b in range(dets.shape[0]):
        n = int(counts[b])
        labels = targets[targets[:, 0] == b, 1:].numpy()
        m = metas[b]
        correct, det = match_image_ref(dets[b, :n].numpy(), labels, m['ratio'][0], m['pad'], m['shape0'], iouv)
        assert np.array_equal(correct.astype(np.uint8), g['correct'][b, :n]), b
        assert det.sum() == correct[:, 0].sum()       # every claimed target has exactly one prediction at IoU 0.5
    assert g['correct'].sum() > 50                    # the fixture exercises the matcher


def test_box_iou_matches_reference(gold):
    g = gold('eval_match.npz')
    dets, counts, targets, _ = eval_case(31, img1=(384, 640))
    t = targets[targets[:, 0] == 0, 2:6].numpy()
    half = t[:, 2:] / np.float32(2)
    tb = np.concatenate((t[:, :2] - half, t[:, :2] + half), 1)
    got = box_iou_ref(dets[0, :int(counts[0]), :4].numpy(), tb)
    assert np.array_equal(got, g['box_iou'])


def test_box_iou_known_answers():
    a = np.array([[0, 0, 4, 1], [0, 0, 2, 2]], np.float32)
    b = np.array([[1, 0, 5, 1], [0, 0, 2, 2], [3, 3, 4, 4]], np.float32)
    k = box_iou_ref(a, b)
    assert k[0, 0] == np.float32(3) / np.float32(5)
    assert k[1, 1] == 1 and k[1, 2] == 0
