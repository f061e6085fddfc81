"""ORACLE (test infrastructure, NOT product code): CPU restatement of the reference's test-time augmentation and
ensemble callers.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it.

Parity status: PINNED -- tests/test_oracle_tta.py checks ``scale_img_ref`` bit for bit against torch's own
F.interpolate + F.pad (what utils/torch_utils.py:247-257 calls) and ``tta_forward_ref`` / ``ensemble_forward_ref``
against outputs of the real reference (Model.forward(augment=True), models/yolo.py:582-597; Ensemble.forward,
models/experimental.py:74-81) stored in tests/golden/tta_*.npz by oracle/make_golden_tta.py.
"""
import math

import numpy as np
import torch

from .ref_forward import ref_forward

F32 = np.float32


def _fma(a, b, c):
    """fp32 fused multiply-add (ATen's vectorised CPU kernel contracts these); exact product in fp64."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(F32)


def _axis(in_size, out_size):
    """ATen area_pixel_compute_source_index + guard_index_and_lambda (align_corners=False, linear)."""
    scale = F32(in_size) / F32(out_size)
    d = np.arange(out_size, dtype=F32)
    src = _fma(scale, d + F32(0.5), F32(-0.5))
    src = np.maximum(src, F32(0))
    i0 = np.minimum(np.floor(src).astype(np.int64), in_size - 1)
    l1 = np.clip(src - i0.astype(F32), F32(0), F32(1)).astype(F32)
    i1 = i0 + (i0 < in_size - 1)
    return i0, i1, (F32(1) - l1).astype(F32), l1


def scale_img_ref(img, ratio=1.0, same_shape=False, gs=32, flip=None):
    """utils/torch_utils.py:247-257 on a (B, C, H, W) float32 array/tensor; ``flip`` (2 | 3 | None) is applied first
    like ``x.flip(fi)`` at models/yolo.py:588.  Returns a float32 torch tensor."""
    x = img.numpy() if isinstance(img, torch.Tensor) else np.asarray(img)
    x = x.astype(F32, copy=False)
    if flip is not None:
        x = np.flip(x, axis=flip)
    if ratio == 1.0:
        return torch.from_numpy(np.ascontiguousarray(x))
    h, w = x.shape[2:]
    s = (int(h * ratio), int(w * ratio))
    y0, y1, ly0, ly1 = _axis(h, s[0])
    x0, x1, lx0, lx1 = _axis(w, s[1])
    a, b = x[:, :, y0][:, :, :, x0], x[:, :, y0][:, :, :, x1]
    c, d = x[:, :, y1][:, :, :, x0], x[:, :, y1][:, :, :, x1]
    top = _fma(a, lx0, (b * lx1).astype(F32))
    bot = _fma(c, lx0, (d * lx1).astype(F32))
    out = _fma(top, ly0[:, None], (bot * ly1[:, None]).astype(F32))
    if not same_shape:
        h, w = [math.ceil(v * ratio / gs) * gs for v in (h, w)]
    full = np.full(x.shape[:2] + (h, w), F32(0.447), dtype=F32)
    full[:, :, :s[0], :s[1]] = out
    return torch.from_numpy(full)


def tta_forward_ref(cfg, sd, x, fused=True, gs=None):
    """Model.forward(x, augment=True): scales 1 / 0.83 / 0.67, the 0.83 pass left-right flipped; returns pred only."""
    img_size = x.shape[-2:]
    if gs is None:      # int(model.stride.max()): 32 for three-level heads, 64 for the P6 four-level heads
        gs = 8 * 2 ** (len(cfg['anchors']) - 1)
    ys = []
    for si, fi in zip((1, 0.83, 0.67), (None, 3, None)):
        xi = scale_img_ref(x, si, gs=gs, flip=fi)
        yi = ref_forward(cfg, sd, xi, fused=fused)[0].clone()
        yi[..., :4] /= si
        if fi == 2:
            yi[..., 1] = img_size[0] - yi[..., 1]
        elif fi == 3:
            yi[..., 0] = img_size[1] - yi[..., 0]
        ys.append(yi)
    return torch.cat(ys, 1)


def ensemble_forward_ref(members, x, augment=False, gs=None):
    """Ensemble.forward: members = [(cfg, state_dict), ...]; NMS-ensemble = concat along dim 1."""
    ys = []
    for cfg, sd in members:
        ys.append(tta_forward_ref(cfg, sd, x, gs=gs) if augment else ref_forward(cfg, sd, x, fused=True)[0])
    return torch.cat(ys, 1)
