"""Evaluation-side callers of the hot path, on the GPU.

* ``box_iou(box1, box2)`` -- drop-in for utils/general.py:464-486 (``yb_box_iou``).
* ``match_batch(out, counts, targets, img1_shape, metas, iouv)`` -- the per-image statistics loop of test.py:128-213
  for the whole batch in one launch (``yb_match_detections``): native-space predictions (test.py:143-144), native-space
  label boxes (test.py:187-190), per-class best IoU, greedy target claim in confidence order, ``correct`` (n, niou).
  Bit-identical to the reference.
* ``NMS`` / ``autoShape`` / ``Detections`` -- the wrappers of models/common.py:852-932 over this package's kernels
  (letterbox, Model, non_max_suppression, scale_coords).  File/URI/PIL loading, plotting and pandas export are host
  conveniences outside the path and are not provided.
"""
import ctypes as C

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .nms import non_max_suppression
from .preprocess import _gain_pad, letterbox_batch, rescale_detections


def _f32_cuda(t, what):
    if not isinstance(t, torch.Tensor) or t.device.type != 'cuda':
        raise RuntimeError(f'yolo_series_b200.{what} needs CUDA tensors (no CPU fallback)')
    return t if t.dtype == torch.float32 else t.float()


def box_iou(box1, box2):
    """Return intersection-over-union (Jaccard index) of boxes, (N, 4) x (M, 4) xyxy -> (N, M)."""
    b1, b2 = _f32_cuda(box1, 'box_iou'), _f32_cuda(box2, 'box_iou')
    if b1.dim() != 2 or b2.dim() != 2 or b1.shape[1] != 4 or b2.shape[1] != 4:
        raise ValueError('box_iou expects (N, 4) and (M, 4) tensors')
    if b1.stride(1) != 1:
        b1 = b1.contiguous()
    if b2.stride(1) != 1:
        b2 = b2.contiguous()
    n, m = b1.shape[0], b2.shape[0]
    out = torch.empty((n, m), dtype=torch.float32, device=b1.device)
    with torch.cuda.device(b1.device):
        _lib.check(_lib.lib().yb_box_iou(C.c_void_p(b1.data_ptr()), n, b1.stride(0) if n > 1 else 4,
                                         C.c_void_p(b2.data_ptr()), m, b2.stride(0) if m > 1 else 4,
                                         C.c_void_p(out.data_ptr()), _lib.stream_ptr()), 'yb_box_iou')
    return out if box1.dtype == torch.float32 else out.to(box1.dtype)


def match_batch(out, counts, targets, img1_shape, metas, iouv, use_ratio_pad=True, return_detected=False):
    """``correct`` of test.py:180-208 for every image of a batch.

    out (B, max_det, 6) fp32 + counts (B,) int32: packed NMS output in inference space (``Model.detect(packed=True)``
    or ``pack_detections`` of a list).  targets (T, 6) rows [image, class, x, y, w, h] in inference-space pixels
    (test.py:109), grouped by image in ascending order (collate_fn).  ``metas[i]`` = {'shape0', 'ratio', 'pad'} as
    ``letterbox_batch`` returns them (= ``shapes[si]`` of test.py:144).  Returns ``correct`` (B, max_det, niou) bool
    on the device (rows beyond counts[b] are False); with ``return_detected`` also the (T,) bool claim mask."""
    out = _f32_cuda(out, 'match_batch').contiguous()
    B, max_det = out.shape[0], out.shape[1]
    dev = out.device
    targets = torch.as_tensor(targets, dtype=torch.float32).to(dev, non_blocking=True).contiguous().view(-1, 6)
    iouv = torch.as_tensor(iouv, dtype=torch.float32).to(dev, non_blocking=True).contiguous()
    counts = counts.to(device=dev, dtype=torch.int32).contiguous()
    table = []
    for m in metas:
        rp = ((m['ratio'][0], m['ratio'][1]), m['pad']) if use_ratio_pad else None
        gain, px, py = _gain_pad(img1_shape, m['shape0'], rp)
        table.append([gain, px, py, float(m['shape0'][1]), float(m['shape0'][0])])
    tab = torch.tensor(table, dtype=torch.float32).to(dev, non_blocking=True)
    niou, T = iouv.numel(), targets.shape[0]
    correct = torch.empty((B, max_det, niou), dtype=torch.uint8, device=dev)
    detected = torch.zeros(max(T, 1), dtype=torch.uint8, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().yb_match_detections(C.c_void_p(out.data_ptr()), C.c_void_p(counts.data_ptr()), B, max_det,
                                                  C.c_void_p(targets.data_ptr()) if T else None, T,
                                                  C.c_void_p(tab.data_ptr()), C.c_void_p(iouv.data_ptr()), niou,
                                                  C.c_void_p(correct.data_ptr()), C.c_void_p(detected.data_ptr()),
                                                  C.c_void_p(status.data_ptr()), _lib.stream_ptr()), 'yb_match_detections')
    correct = correct.bool()
    correct._yb_status = status           # read with check_match_status() after the caller's own synchronisation
    return (correct, detected[:T].bool()) if return_detected else correct


def check_match_status(correct):
    """Raises if the targets handed to ``match_batch`` were not grouped by image (one 4-byte read-back)."""
    st = getattr(correct, '_yb_status', None)
    if st is not None and int(st.item()) != 0:
        raise ValueError('match_batch: targets must be grouped by image index in ascending order with indices in '
                         '[0, B) (utils/datasets.py:621-626 collate_fn order)')


def pack_detections(dets, max_det=None):
    """list of (n_i, 6) tensors (``non_max_suppression`` output) -> (out (B, max_det, 6), counts (B,) int32)."""
    B = len(dets)
    n = [int(d.shape[0]) for d in dets]
    max_det = max_det or max(max(n), 1)
    dev = dets[0].device
    out = torch.zeros((B, max_det, 6), dtype=torch.float32, device=dev)
    for i, d in enumerate(dets):
        out[i, :n[i]] = d[:, :6]
    return out, torch.tensor(n, dtype=torch.int32).to(dev, non_blocking=True)


class NMS(nn.Module):
    """Non-Maximum Suppression (NMS) module (models/common.py:852-862)."""
    conf = 0.25
    iou = 0.45
    classes = None

    def forward(self, x):
        return non_max_suppression(x[0], conf_thres=self.conf, iou_thres=self.iou, classes=self.classes)


class Detections:
    """Result holder of ``autoShape`` (models/common.py:935-958): pixel / normalised xyxy and xywh lists."""

    def __init__(self, imgs, pred, files, names=None, shape=None):
        self.imgs, self.pred, self.files, self.names = imgs, pred, files, names
        gn = [torch.tensor([im.shape[1], im.shape[0], im.shape[1], im.shape[0], 1., 1.], device=pred[0].device)
              for im in imgs]
        self.xyxy = pred
        self.xywh = [torch.cat(((p[:, :2] + p[:, 2:4]) / 2, p[:, 2:4] - p[:, :2], p[:, 4:]), 1) for p in pred]
        self.xyxyn = [x / g for x, g in zip(self.xyxy, gn)]
        self.xywhn = [x / g for x, g in zip(self.xywh, gn)]
        self.n = len(pred)
        self.s = shape

    def __len__(self):
        return self.n

    def tolist(self):
        return [Detections([self.imgs[i]], [self.pred[i]], [self.files[i]], self.names, self.s) for i in range(self.n)]


class autoShape(nn.Module):
    """Input-robust wrapper (models/common.py:865-932): HWC RGB uint8 frames (numpy arrays or tensors, any sizes) ->
    letterbox to a common stride-multiple shape -> model -> NMS -> boxes in each frame's own pixels.  All four steps
    run on the device (yb_letterbox, engine, yb_decode_nms, yb_scale_coords)."""
    conf = 0.25
    iou = 0.45
    classes = None

    def __init__(self, model):
        super().__init__()
        self.model = model.eval()
        self.stride = model.stride
        self.names = model.names

    def autoshape(self):
        return self

    @torch.no_grad()
    def forward(self, imgs, size=640, augment=False, profile=False):
        if isinstance(imgs, torch.Tensor) and imgs.dim() == 4:      # BCHW batch: straight to the model (common.py:892-894)
            return self.model(imgs, augment, profile)
        imgs = list(imgs) if isinstance(imgs, (list, tuple)) else [imgs]
        frames, shape0, shape1 = [], [], []
        for im in imgs:
            if isinstance(im, torch.Tensor):
                im = im.cpu().numpy() if im.device.type != 'cuda' else im
            if im.shape[0] < 5:                                    # CHW -> HWC (common.py:907-908)
                im = im.permute(1, 2, 0) if isinstance(im, torch.Tensor) else im.transpose((1, 2, 0))
            if im.ndim == 2:
                im = im[:, :, None].repeat(1, 1, 3) if isinstance(im, torch.Tensor) else np.tile(im[:, :, None], 3)
            im = im[:, :, :3]
            s = tuple(int(v) for v in im.shape[:2])
            shape0.append(s)
            g = size / max(s)
            shape1.append([y * g for y in s])
            frames.append(im if isinstance(im, torch.Tensor) else np.ascontiguousarray(im))
        gs = int(self.stride.max())
        shape1 = [int(np.ceil(v / gs) * gs) for v in np.stack(shape1, 0).max(0)]   # make_divisible
        dev = next(self.model.parameters()).device
        x, metas = letterbox_batch(frames, new_shape=shape1, auto=False, swap_rb=False, device=dev)
        if augment:
            y = self.model.detect(x, self.conf, self.iou, classes=self.classes, augment=True, packed=True)
        else:
            y = self.model.detect(x, self.conf, self.iou, classes=self.classes, packed=True)
        out, meta = y
        rescale_detections(out, meta[:-1], shape1, metas, round_=False)       # scale_coords(shape1, y[i][:, :4], shape0[i])
        host = meta.cpu()
        if int(host[-1]) != 0:
            raise RuntimeError('yb_nms: candidate capacity exceeded (YB_ERR_OVERFLOW)')
        pred = [out[i, :int(host[i])] for i in range(len(frames))]
        files = [f'image{i}.jpg' for i in range(len(frames))]
        return Detections(frames, pred, files, self.names, tuple(x.shape))
