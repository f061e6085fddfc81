"""ORACLE (test infrastructure, NOT product code): CPU restatement of the evaluation-side matching of the reference
(SURVEY.md 8(f) row 4): box_iou (utils/general.py:464-486) and the per-image prediction -> target assignment of
test.py:179-210.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it.

Parity status: PINNED -- tests/test_oracle_eval.py checks ``box_iou_ref`` and ``match_image_ref`` against outputs of
the real reference functions (utils.general.box_iou / scale_coords / xywh2xyxy imported from /root/reference and the
loop of test.py:186-208 executed verbatim) stored in tests/golden/eval_match.npz by oracle/make_golden_tta.py.
"""
import numpy as np
import torch


def box_iou_ref(box1, box2):
    """general.py:464-486 in numpy fp32: (N, 4) x (M, 4) -> (N, M)."""
    b1, b2 = np.asarray(box1, np.float32), np.asarray(box2, np.float32)
    a1 = (b1[:, 2] - b1[:, 0]) * (b1[:, 3] - b1[:, 1])
    a2 = (b2[:, 2] - b2[:, 0]) * (b2[:, 3] - b2[:, 1])
    wh = np.minimum(b1[:, None, 2:4], b2[None, :, 2:4]) - np.maximum(b1[:, None, :2], b2[None, :, :2])
    wh = np.maximum(wh, np.float32(0))
    inter = wh[..., 0] * wh[..., 1]
    with np.errstate(invalid='ignore', divide='ignore'):
        return inter / ((a1[:, None] + a2[None, :]) - inter)


def _scale_coords(coords, gain, pad, shape0):
    """general.py:340-361 with ratio_pad given (test.py:144), fp32, out of place."""
    c = np.array(coords, np.float32, copy=True)
    c[:, [0, 2]] -= np.float32(pad[0])
    c[:, [1, 3]] -= np.float32(pad[1])
    c[:, :4] /= np.float32(gain)
    c[:, [0, 2]] = np.clip(c[:, [0, 2]], 0, np.float32(shape0[1]))
    c[:, [1, 3]] = np.clip(c[:, [1, 3]], 0, np.float32(shape0[0]))
    return c


def match_image_ref(pred, labels, gain, pad, shape0, iouv):
    """test.py:179-208 for one image.  pred (n, 6) inference-space NMS rows; labels (nl, 5) [cls, x, y, w, h] pixels,
    inference space; iouv (niou,) fp32.  Returns (correct (n, niou) bool, detected (nl,) bool)."""
    pred = np.asarray(pred, np.float32)
    labels = np.asarray(labels, np.float32).reshape(-1, 5)
    iouv = np.asarray(iouv, np.float32)
    n, nl = pred.shape[0], labels.shape[0]
    correct = np.zeros((n, iouv.shape[0]), bool)
    det = np.zeros(nl, bool)
    if n == 0 or nl == 0:
        return correct, det
    predn = _scale_coords(pred[:, :4], gain, pad, shape0)
    xy, half = labels[:, 1:3], labels[:, 3:5] / np.float32(2)
    tbox = _scale_coords(np.concatenate((xy - half, xy + half), 1), gain, pad, shape0)
    tcls = labels[:, 0]
    detected = []
    for cls in np.unique(tcls):
        ti = np.nonzero(cls == tcls)[0]
        pi = np.nonzero(cls == pred[:, 5])[0]
        if pi.shape[0]:
            m = torch.from_numpy(box_iou_ref(predn[pi], tbox[ti]))
            ious, i = m.max(1)            # torch semantics: first maximum, NaN propagates
            ious, i = ious.numpy(), i.numpy()
            seen = set()
            for j in np.nonzero(ious > iouv[0])[0]:
                d = int(ti[i[j]])
                if d not in seen:
                    seen.add(d)
                    detected.append(d)
                    correct[pi[j]] = ious[j] > iouv
                    if len(detected) == nl:
                        break
    det[detected] = True
    return correct, det
