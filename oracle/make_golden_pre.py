"""Generates tests/golden/pre_*.npz by running the REAL reference functions (utils/datasets.py letterbox with the
installed cv2, utils/general.py scale_coords) in the build container:

    python -m oracle.make_golden_pre

Inputs are regenerated from seeds by ``synth_frame`` / ``synth_boxes`` below on both sides; the fixtures hold the
reference outputs (full arrays for the small cases, SHA-256 + a strided sample for the camera-sized ones).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, 'tests', 'golden')

# (tag, src_h, src_w, seed, kwargs of letterbox)
LETTERBOX_CASES = [
    ('down_auto', 90, 120, 1, dict(new_shape=64, auto=True)),
    ('down_fixed', 90, 120, 2, dict(new_shape=(64, 64), auto=False)),
    ('up', 37, 53, 3, dict(new_shape=96, auto=False)),
    ('noscaleup', 37, 53, 4, dict(new_shape=96, auto=False, scaleup=False)),
    ('half', 128, 96, 5, dict(new_shape=64, auto=False)),                 # exact 2x reduction (box average)
    ('fill', 50, 81, 6, dict(new_shape=(64, 96), auto=False, scaleFill=True)),
    ('same', 64, 64, 7, dict(new_shape=64, auto=False)),                  # no resize, no border
    ('tall', 121, 45, 8, dict(new_shape=(96, 64), auto=True, stride=32)),
    ('odd_color', 75, 101, 9, dict(new_shape=(80, 112), auto=False, color=(1, 114, 250))),
    ('vga', 480, 640, 10, dict(new_shape=640, auto=False)),
    ('hd', 720, 1280, 11, dict(new_shape=640, auto=False)),
    ('fullhd_auto', 1080, 1920, 12, dict(new_shape=640, auto=True)),
    ('uxga', 1200, 1600, 13, dict(new_shape=640, auto=False)),
    ('coco_up', 333, 500, 14, dict(new_shape=640, auto=False)),
]

# (tag, img1_shape, img0_shape, n, seed, with ratio_pad from letterbox?)
SCALE_CASES = [
    ('vga', (640, 640), (480, 640), 300, 21, False),
    ('hd_ratio_pad', (640, 640), (720, 1280), 300, 22, True),
    ('rect', (384, 640), (1080, 1920), 257, 23, False),
    ('up', (640, 640), (375, 500), 64, 24, True),
    ('one', (640, 640), (427, 640), 1, 25, False),
]


def synth_frame(h, w, seed):
    """Smooth gradients + texture + noise, uint8 BGR (keeps edges and flat areas in the picture)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.empty((h, w, 3), np.float64)
    for c in range(3):
        fx, fy, ph = rng.uniform(0.01, 0.3), rng.uniform(0.01, 0.3), rng.uniform(0, 6.28)
        img[..., c] = 127.5 + 90 * np.sin(xx * fx + ph) * np.cos(yy * fy) + rng.normal(0, 25, (h, w))
    img[h // 3: h // 3 + max(h // 8, 1), :, :] = 255      # saturated bars exercise the clamps
    img[:, w // 2: w // 2 + max(w // 16, 1), :] = 0
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_boxes(n, img1_shape, seed):
    """xyxy + conf + cls rows in letterboxed coordinates, some outside the image, some on .5 boundaries."""
    rng = np.random.default_rng(seed)
    h, w = img1_shape
    xy = rng.uniform(-20, [w + 20, h + 20], (n, 2))
    wh = rng.uniform(1, 200, (n, 2))
    d = np.concatenate([xy - wh / 2, xy + wh / 2, rng.uniform(0, 1, (n, 1)), rng.integers(0, 80, (n, 1))], 1)
    d = d.astype(np.float32)
    d[::7, :4] = np.rint(d[::7, :4] * 2) / 2
    return d


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    sys.path.insert(0, ROOT)
    from oracle.make_golden import import_reference
    import_reference()
    import torch
    from utils.datasets import letterbox as ref_letterbox
    from utils.general import scale_coords as ref_scale_coords
    os.makedirs(GOLD, exist_ok=True)
    out = {}
    for tag, h, w, seed, kw in LETTERBOX_CASES:
        img, ratio, pad = ref_letterbox(synth_frame(h, w, seed), **kw)
        out[f'lb_{tag}_shape'] = np.array(img.shape)
        out[f'lb_{tag}_ratio'] = np.array(ratio, np.float64)
        out[f'lb_{tag}_pad'] = np.array(pad, np.float64)
        if img.size <= 64 * 1024:
            out[f'lb_{tag}_img'] = img
        else:
            out[f'lb_{tag}_sha'] = np.frombuffer(bytes.fromhex(digest(img)), np.uint8)
            out[f'lb_{tag}_sub'] = img[::17, ::13].copy()
        print('letterbox', tag, img.shape, ratio, pad)
    for tag, s1, s0, n, seed, use_rp in SCALE_CASES:
        d = synth_boxes(n, s1, seed)
        rp = None
        if use_rp:
            _, ratio, pad = ref_letterbox(np.zeros((s0[0], s0[1], 3), np.uint8), s1, auto=False)
            rp = (ratio, pad)
        t = torch.from_numpy(d.copy())
        ref_scale_coords(s1, t[:, :4], s0, rp)
        out[f'sc_{tag}'] = t.numpy().copy()
        out[f'sc_{tag}_round'] = t[:, :4].round().numpy().copy()
        print('scale_coords', tag, t[:2, :4].tolist())
    np.savez_compressed(os.path.join(GOLD, 'pre_post.npz'), **out)


if __name__ == '__main__':
    main()
