"""Pre- and post-processing either side of the hot path, on the GPU.

Drop-ins for the reference functions, same names / argument order / defaults / return values:

* ``letterbox(img, new_shape=(640, 640), color=(114, 114, 114), auto=True, scaleFill=False, scaleup=True,
  stride=32)`` -- utils/datasets.py:1277-1307.  ``img`` is an (H, W, 3) uint8 frame (numpy array or torch
  tensor; a host array is copied to the current CUDA device).  The resize has the arithmetic of
  ``cv2.resize(..., INTER_LINEAR)``, so the pixels are the reference's bit for bit.
* ``scale_coords(img1_shape, coords, img0_shape, ratio_pad=None)`` and ``clip_coords(boxes, img_shape)`` --
  utils/general.py:340-361, in place on a CUDA tensor.

Batched forms that keep the whole detect.py loop body (datasets.py:195-199 -> detect.py:100-104 -> model -> NMS ->
detect.py:183) on the device:

* ``letterbox_batch(frames, ...)`` -> ``(x (B, 3, H, W) uint8 RGB, metas)``: letterbox + BGR->RGB + HWC->CHW in one
  kernel launch for up to 32 frames of arbitrary sizes; ``x`` is what ``Model`` takes (the first-layer kernel applies
  the /255).
* ``rescale_detections(out, counts, img1_shape, metas, round_=True)``: scale_coords on the packed NMS output.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


def _geometry(shape, new_shape, auto, scaleFill, scaleup, stride):
    """Sizes and offsets of datasets.py:1283-1304 for a source of ``shape`` (h, w)."""
    if isinstance(new_shape, int):
        new_shape = (new_shape, new_shape)
    h0, w0 = int(shape[0]), int(shape[1])
    r = min(new_shape[0] / h0, new_shape[1] / w0)
    if not scaleup:
        r = min(r, 1.0)
    ratio = (r, r)
    new_w, new_h = int(round(w0 * r)), int(round(h0 * r))
    dw, dh = new_shape[1] - new_w, new_shape[0] - new_h
    if auto:
        dw, dh = dw % stride, dh % stride
    elif scaleFill:
        dw, dh = 0.0, 0.0
        new_w, new_h = new_shape[1], new_shape[0]
        ratio = (new_shape[1] / w0, new_shape[0] / h0)
    dw, dh = dw / 2, dh / 2
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    return dict(new_w=new_w, new_h=new_h, top=top, left=left, out_h=new_h + top + bottom,
                out_w=new_w + left + right, ratio=ratio, pad=(dw, dh), shape0=(h0, w0))


def _as_cuda_frame(img, device):
    if isinstance(img, np.ndarray):
        img = torch.from_numpy(np.ascontiguousarray(img))
    if img.dtype != torch.uint8 or img.dim() != 3 or img.shape[2] != 3:
        raise ValueError(f'letterbox expects an (H, W, 3) uint8 frame, got {tuple(img.shape)} {img.dtype}')
    if img.device.type != 'cuda':
        if device is None:
            if not torch.cuda.is_available():
                raise RuntimeError('yolo_series_b200.letterbox needs a CUDA device (no CPU fallback)')
            device = torch.device('cuda', torch.cuda.current_device())
        img = img.to(device, non_blocking=True)
    if img.stride(2) != 1 or img.stride(1) != 3:
        img = img.contiguous()
    return img


def _launch(frames, geos, out, layout, swap_rb, color):
    B = len(frames)
    srcs = (_lib.LetterboxSrc * B)()
    for i, (f, g) in enumerate(zip(frames, geos)):
        s = srcs[i]
        s.src, s.src_pitch = f.data_ptr(), f.stride(0)
        s.src_h, s.src_w = f.shape[0], f.shape[1]
        s.new_w, s.new_h, s.top, s.left = g['new_w'], g['new_h'], g['top'], g['left']
    border = (C.c_uint8 * 3)(*[int(c) & 255 for c in color])
    H, W = (out.shape[2], out.shape[3]) if layout == _lib.LAYOUT_CHW else (out.shape[1], out.shape[2])
    with torch.cuda.device(out.device):       # launches follow the tensors' device, like torch ops
        _lib.check(_lib.lib().yb_letterbox(srcs, B, C.c_void_p(out.data_ptr()), H, W, layout, int(bool(swap_rb)), border,
                                           _lib.stream_ptr()), 'yb_letterbox')


def letterbox(img, new_shape=(640, 640), color=(114, 114, 114), auto=True, scaleFill=False, scaleup=True, stride=32):
    """Resize and pad image while meeting stride-multiple constraints.  Returns (img, ratio, (dw, dh))."""
    f = _as_cuda_frame(img, None)
    g = _geometry(f.shape[:2], new_shape, auto, scaleFill, scaleup, stride)
    out = torch.empty((1, g['out_h'], g['out_w'], 3), dtype=torch.uint8, device=f.device)
    _launch([f], [g], out, _lib.LAYOUT_HWC, False, color)
    return out[0], g['ratio'], g['pad']


def letterbox_batch(frames, new_shape=(640, 640), color=(114, 114, 114), auto=False, scaleFill=False, scaleup=True,
                    stride=32, swap_rb=True, out=None, device=None):
    """frames: sequence of (H_i, W_i, 3) uint8 BGR frames (sizes may differ).  Returns ``(x, metas)``: ``x`` the
    (B, 3, H, W) uint8 RGB batch on the device, ``metas[i] = {'ratio', 'pad', 'shape0'}`` = what the reference keeps
    per image as ``shapes`` / ``ratio_pad`` (datasets.py:617, test.py:144)."""
    frames = [_as_cuda_frame(f, device) for f in frames]
    if not frames:
        raise ValueError('letterbox_batch: no frames')
    geos = [_geometry(f.shape[:2], new_shape, auto, scaleFill, scaleup, stride) for f in frames]
    H, W = geos[0]['out_h'], geos[0]['out_w']
    if any((g['out_h'], g['out_w']) != (H, W) for g in geos):
        raise ValueError('letterbox_batch: auto=True gave different padded shapes; batch frames of one aspect '
                         'ratio or use auto=False')
    if out is None:
        out = torch.empty((len(frames), 3, H, W), dtype=torch.uint8, device=frames[0].device)
    elif tuple(out.shape) != (len(frames), 3, H, W) or out.dtype != torch.uint8 or not out.is_contiguous():
        raise ValueError(f'letterbox_batch: out must be a contiguous uint8 ({len(frames)}, 3, {H}, {W}) tensor')
    _launch(frames, geos, out, _lib.LAYOUT_CHW, swap_rb, color)
    return out, [dict(ratio=g['ratio'], pad=g['pad'], shape0=g['shape0']) for g in geos]


# ---------------------------------------------------------------------------------------------- after NMS
def _gain_pad(img1_shape, img0_shape, ratio_pad):
    if ratio_pad is None:   # general.py:342-344
        gain = min(img1_shape[0] / img0_shape[0], img1_shape[1] / img0_shape[1])
        pad = (img1_shape[1] - img0_shape[1] * gain) / 2, (img1_shape[0] - img0_shape[0] * gain) / 2
    else:
        gain, pad = ratio_pad[0][0], ratio_pad[1]
    return float(gain), float(pad[0]), float(pad[1])


def _scale(dets, B, max_rows, counts, table, do_round):
    if dets.dtype != torch.float32 or dets.device.type != 'cuda' or dets.stride(-1) != 1:
        raise ValueError('scale_coords needs a float32 CUDA tensor with unit column stride (no CPU fallback)')
    tab = torch.tensor(table, dtype=torch.float32).to(dets.device, non_blocking=True)
    row_stride = dets.stride(-2)
    cptr = C.c_void_p(counts.data_ptr()) if counts is not None else None
    with torch.cuda.device(dets.device):
        _lib.check(_lib.lib().yb_scale_coords(C.c_void_p(dets.data_ptr()), B, max_rows, row_stride, cptr,
                                              C.c_void_p(tab.data_ptr()), int(bool(do_round)), _lib.stream_ptr()),
                   'yb_scale_coords')


def scale_coords(img1_shape, coords, img0_shape, ratio_pad=None):
    """Rescale coords (xyxy) from img1_shape to img0_shape, in place; returns ``coords``."""
    n = coords.shape[0]
    if n == 0:
        return coords
    if coords.dim() != 2 or coords.shape[1] < 4 or (n > 1 and coords.stride(0) < 4):
        raise ValueError('scale_coords expects an (n, >=4) tensor')
    gain, px, py = _gain_pad(img1_shape, img0_shape, ratio_pad)
    _scale(coords, 1, n, None, [[gain, px, py, float(img0_shape[1]), float(img0_shape[0])]], False)
    return coords


def clip_coords(boxes, img_shape):
    """Clip xyxy boxes to the image (height, width), in place (general.py:356-361)."""
    if boxes.shape[0]:
        _scale(boxes, 1, boxes.shape[0], None, [[1.0, 0.0, 0.0, float(img_shape[1]), float(img_shape[0])]], False)


def rescale_detections(out, counts, img1_shape, metas, round_=True, use_ratio_pad=False):
    """scale_coords (+ detect.py:183 ``.round()``) on the packed NMS output ``out`` (B, max_det, 6) for all images in
    one launch, in place.  ``metas[i]['shape0']`` is image i's original (h, w); with ``use_ratio_pad`` the letterbox's
    own ``ratio`` / ``pad`` are used (test.py:144 form) instead of recomputing them (detect.py:183 form)."""
    B, max_det = out.shape[0], out.shape[1]
    table = []
    for m in metas:
        rp = ((m['ratio'][0], m['ratio'][1]), m['pad']) if use_ratio_pad else None
        gain, px, py = _gain_pad(img1_shape, m['shape0'], rp)
        table.append([gain, px, py, float(m['shape0'][1]), float(m['shape0'][0])])
    _scale(out, B, max_det, counts, table, round_)
    return out
