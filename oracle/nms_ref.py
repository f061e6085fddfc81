"""ORACLE (test infrastructure, NOT product code): CPU restatement of the reference's
``non_max_suppression`` (utils/general.py:628-720) with ``max_det`` / ``max_nms`` as parameters
(hard-coded 300 / 30000 in the reference, lines 641-642) and without the 10 s ``time_limit`` abort
(lines 643, 716-718).  The greedy step (general.py:704, torchvision.ops.nms) is the C restatement in
oracle/nms_greedy.c.

Parity status: PINNED -- tests/test_oracle_nms.py checks (a) the greedy step against
torchvision.ops.nms and its known-answer vectors, (b) this whole function against the
real reference function imported from /root/reference (golden fixtures in tests/golden/, generated
by oracle/make_golden.py) at max_det=300, max_nms=30000.
"""
import ctypes as C

import numpy as np
import torch

from . import build as _build

_lib = None


def _greedy_lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_build.build())
        _lib.yb_oracle_nms_sorted.restype = C.c_int64
        _lib.yb_oracle_nms_sorted.argtypes = [C.c_void_p, C.c_int64, C.c_double, C.c_int64, C.c_void_p]
    return _lib


def greedy_nms(boxes, scores, iou_thres, max_keep=0):
    """torchvision.ops.nms semantics on CPU tensors; returns kept indices (int64) in score order."""
    n = boxes.shape[0]
    if n == 0:
        return torch.zeros((0,), dtype=torch.int64)
    order = torch.sort(scores.float(), stable=True, descending=True)[1]      # nms_kernel.cpp: stable sort
    sb = np.ascontiguousarray(boxes.float()[order].numpy(), dtype=np.float32)
    keep = np.empty(n, dtype=np.int64)
    nk = _greedy_lib().yb_oracle_nms_sorted(sb.ctypes.data, n, float(iou_thres), int(max_keep), keep.ctypes.data)
    return order[torch.from_numpy(keep[:nk])]


def xywh2xyxy(x):  # general.py:275-282
    y = x.clone()
    y[:, 0] = x[:, 0] - x[:, 2] / 2
    y[:, 1] = x[:, 1] - x[:, 3] / 2
    y[:, 2] = x[:, 0] + x[:, 2] / 2
    y[:, 3] = x[:, 1] + x[:, 3] / 2
    return y


@torch.no_grad()
def non_max_suppression_ref(prediction, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False,
                            multi_label=False, labels=(), max_det=300, max_nms=30000, stable_cut=True,
                            early_exit=True):
    """``early_exit=False`` makes the greedy step visit every candidate like torchvision.ops.nms does (same result; used
    when this function is TIMED as the reference's CPU cost -- stopping at max_det is a shortcut the reference lacks)."""
    prediction = prediction.float().cpu()
    nc = prediction.shape[2] - 5                                   # :636
    xc = prediction[..., 4] > conf_thres                           # :637
    max_wh = 4096                                                  # :640
    multi_label &= nc > 1                                          # :645
    output = [torch.zeros((0, 6)) for _ in range(prediction.shape[0])]   # :649 (distinct tensors)
    for xi, x in enumerate(prediction):                            # :650
        x = x[xc[xi]]                                              # :653
        if labels and len(labels[xi]):                             # :656-662
            l = torch.as_tensor(labels[xi], dtype=torch.float32)
            v = torch.zeros((len(l), nc + 5))
            v[:, :4] = l[:, 1:5]
            v[:, 4] = 1.0
            v[range(len(l)), l[:, 0].long() + 5] = 1.0
            x = torch.cat((x, v), 0)
        if not x.shape[0]:                                         # :665
            continue
        if nc == 1:                                                # :669-673
            x[:, 5:] = x[:, 4:5]
        else:
            x[:, 5:] *= x[:, 4:5]
        box = xywh2xyxy(x[:, :4])                                  # :676
        if multi_label:                                            # :679-684
            i, j = (x[:, 5:] > conf_thres).nonzero(as_tuple=False).T
            x = torch.cat((box[i], x[i, j + 5, None], j[:, None].float()), 1)
        else:
            conf, j = x[:, 5:].max(1, keepdim=True)
            x = torch.cat((box, conf, j.float()), 1)[conf.view(-1) > conf_thres]
        if classes is not None:                                    # :687-688
            x = x[(x[:, 5:6] == torch.tensor(classes, dtype=torch.float32)).any(1)]
        n = x.shape[0]                                             # :695-699
        if not n:
            continue
        elif n > max_nms:
            # reference: argsort(descending=True) (unstable); stable here = one legal outcome (A.6)
            order = torch.sort(x[:, 4], stable=stable_cut, descending=True)[1]
            x = x[order[:max_nms]]
        c = x[:, 5:6] * (0 if agnostic else max_wh)                # :702
        boxes, scores = x[:, :4] + c, x[:, 4]                      # :703
        i = greedy_nms(boxes, scores, iou_thres, max_keep=max_det if early_exit else 0)  # :704
        if i.shape[0] > max_det:                                   # :705-706
            i = i[:max_det]
        output[xi] = x[i]                                          # :715
    return output
