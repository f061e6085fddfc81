"""ORACLE (test infrastructure, not product code): CPU restatement of the pre-/post-processing around the hot
path.

* ``letterbox_ref``      -- reference utils/datasets.py:1277-1307 (geometry) + detect-side conversion
                            datasets.py:195-199 (BGR->RGB, HWC->CHW).
* ``resize_linear_u8``   -- the arithmetic of ``cv2.resize(..., interpolation=cv2.INTER_LINEAR)`` on uint8, which the
                            reference calls at datasets.py:1303.  OpenCV is a third-party dependency the reference does
                            not pin (no requirements file); installed here: opencv-python 4.13.  Restated from its
                            published algorithm (imgproc resize: 11-bit fixed-point coefficients, horizontal pass in
                            int32, vertical pass ``((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2``, exact 2x
                            reduction taken as a 2x2 box average) and PINNED against the installed binary by
                            ``oracle/make_golden_pre.py`` / ``tests/test_oracle_pre.py`` (bit-exact on every size tried).
* ``scale_coords_ref``   -- reference utils/general.py:340-361 (+ ``.round()`` of detect.py:183) in fp32 numpy, pinned
                            against the reference function itself (``oracle/make_golden_pre.py``).
"""
import numpy as np

COEF_BITS = 11
COEF_ONE = 1 << COEF_BITS


def _round_half_even(v):
    return np.rint(v)


def _axis_coeffs(dst_n, src_n, clamp_edges):
    """Source index and the two 11-bit weights per destination index (OpenCV: xofs/ialpha, yofs/ibeta)."""
    inv_scale = float(dst_n) / float(src_n)
    scale = 1.0 / inv_scale
    d = np.arange(dst_n, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int64)
    f = (f - s.astype(np.float32)).astype(np.float32)
    if clamp_edges:                       # horizontal pass: edge taps collapse onto the border pixel
        lo = s < 0
        f[lo] = 0.0
        s[lo] = 0
        hi = s >= src_n - 1
        f[hi] = 0.0
        s[hi] = src_n - 1
    w0 = _round_half_even((np.float32(1.0) - f) * np.float32(COEF_ONE))
    w1 = _round_half_even(f * np.float32(COEF_ONE))
    w0 = np.clip(w0, -32768, 32767).astype(np.int64)
    w1 = np.clip(w1, -32768, 32767).astype(np.int64)
    return s, w0, w1


def resize_linear_u8(img, dst_w, dst_h):
    """img (H, W, C) uint8 -> (dst_h, dst_w, C) uint8, as cv2.resize(img, (dst_w, dst_h), INTER_LINEAR)."""
    img = np.asarray(img)
    assert img.dtype == np.uint8 and img.ndim == 3
    sh, sw = img.shape[:2]
    if (sw, sh) == (dst_w, dst_h):
        return img.copy()
    if sw == 2 * dst_w and sh == 2 * dst_h:   # exact 2x reduction: OpenCV switches to the 2x2 box average
        a = img.astype(np.int64)
        s = a[0::2, 0::2] + a[0::2, 1::2] + a[1::2, 0::2] + a[1::2, 1::2]
        return ((s + 2) >> 2).astype(np.uint8)
    sx, a0, a1 = _axis_coeffs(dst_w, sw, True)
    sy, b0, b1 = _axis_coeffs(dst_h, sh, False)
    src = img.astype(np.int64)
    sx1 = np.minimum(sx + 1, sw - 1)
    rows = src[:, sx, :] * a0[None, :, None] + src[:, sx1, :] * a1[None, :, None]     # (sh, dst_w, C) int32 range
    y0 = np.clip(sy, 0, sh - 1)
    y1 = np.clip(sy + 1, 0, sh - 1)
    r0, r1 = rows[y0], rows[y1]
    out = (((b0[:, None, None] * (r0 >> 4)) >> 16) + ((b1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def letterbox_geometry(shape, new_shape=(640, 640), auto=True, scaleFill=False, scaleup=True, stride=32):
    """datasets.py:1279-1304 without the pixels: returns (new_unpad (w, h), (top, bottom, left, right), ratio,
    (dw, dh))."""
    if isinstance(new_shape, int):
        new_shape = (new_shape, new_shape)
    r = min(new_shape[0] / shape[0], new_shape[1] / shape[1])
    if not scaleup:
        r = min(r, 1.0)
    ratio = r, r
    new_unpad = int(round(shape[1] * r)), int(round(shape[0] * r))
    dw, dh = new_shape[1] - new_unpad[0], new_shape[0] - new_unpad[1]
    if auto:
        dw, dh = dw % stride, dh % stride
    elif scaleFill:
        dw, dh = 0.0, 0.0
        new_unpad = (new_shape[1], new_shape[0])
        ratio = new_shape[1] / shape[1], new_shape[0] / shape[0]
    dw /= 2
    dh /= 2
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    return new_unpad, (top, bottom, left, right), ratio, (dw, dh)


def letterbox_ref(img, new_shape=(640, 640), color=(114, 114, 114), auto=True, scaleFill=False, scaleup=True,
                  stride=32):
    """(H, W, 3) uint8 -> (letterboxed (H', W', 3) uint8, ratio, (dw, dh)); datasets.py:1277-1307."""
    shape = img.shape[:2]
    new_unpad, (top, bottom, left, right), ratio, dwdh = letterbox_geometry(shape, new_shape, auto, scaleFill,
                                                                             scaleup, stride)
    if shape[::-1] != new_unpad:
        img = resize_linear_u8(img, new_unpad[0], new_unpad[1])
    h, w = img.shape[:2]
    out = np.empty((h + top + bottom, w + left + right, 3), np.uint8)
    out[...] = np.asarray(color, np.uint8)[None, None, :]
    out[top:top + h, left:left + w] = img
    return out, ratio, dwdh


def to_model_input(img_bgr_hwc):
    """datasets.py:198-199: BGR -> RGB, HWC -> CHW, contiguous uint8."""
    return np.ascontiguousarray(img_bgr_hwc[:, :, ::-1].transpose(2, 0, 1))


def scale_coords_ref(img1_shape, coords, img0_shape, ratio_pad=None, do_round=False):
    """general.py:340-361 on an (n, >=4) float32 array (returns a new array); ``do_round`` adds detect.py:183's
    ``.round()`` (half to even, like torch.round)."""
    c = np.array(coords, dtype=np.float32, copy=True)
    if ratio_pad is None:
        gain = min(img1_shape[0] / img0_shape[0], img1_shape[1] / img0_shape[1])
        pad = (img1_shape[1] - img0_shape[1] * gain) / 2, (img1_shape[0] - img0_shape[0] * gain) / 2
    else:
        gain = ratio_pad[0][0]
        pad = ratio_pad[1]
    px, py, g = np.float32(pad[0]), np.float32(pad[1]), np.float32(gain)
    c[:, [0, 2]] = c[:, [0, 2]] - px
    c[:, [1, 3]] = c[:, [1, 3]] - py
    c[:, :4] = c[:, :4] / g
    w, h = np.float32(img0_shape[1]), np.float32(img0_shape[0])
    c[:, 0] = np.clip(c[:, 0], 0, w)
    c[:, 1] = np.clip(c[:, 1], 0, h)
    c[:, 2] = np.clip(c[:, 2], 0, w)
    c[:, 3] = np.clip(c[:, 3], 0, h)
    if do_round:
        c[:, :4] = np.rint(c[:, :4])
    return c
