"""CPU: the pre-/post-processing oracle (oracle/preprocess_ref.py) against fixtures produced by the REAL reference
(utils/datasets.py letterbox + installed cv2, utils/general.py scale_coords; oracle/make_golden_pre.py)."""
import hashlib
import os

import numpy as np
import pytest

from oracle import preprocess_ref as R
from oracle.make_golden_pre import LETTERBOX_CASES, SCALE_CASES, synth_boxes, synth_frame

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'pre_post.npz'))


def check_letterbox_against_golden(tag, img):
    """Shared with the GPU tests: img is an (H, W, 3) uint8 numpy array."""
    assert list(img.shape) == GOLD[f'lb_{tag}_shape'].tolist()
    if f'lb_{tag}_img' in GOLD:
        assert np.array_equal(img, GOLD[f'lb_{tag}_img'])
    else:
        assert np.array_equal(img[::17, ::13], GOLD[f'lb_{tag}_sub'])
        sha = hashlib.sha256(np.ascontiguousarray(img).tobytes()).digest()
        assert sha == GOLD[f'lb_{tag}_sha'].tobytes()


@pytest.mark.parametrize('case', LETTERBOX_CASES, ids=[c[0] for c in LETTERBOX_CASES])
def test_letterbox_oracle_matches_reference(case):
    tag, h, w, seed, kw = case
    img, ratio, pad = R.letterbox_ref(synth_frame(h, w, seed), **kw)
    check_letterbox_against_golden(tag, img)
    assert np.array_equal(np.array(ratio, np.float64), GOLD[f'lb_{tag}_ratio'])
    assert np.array_equal(np.array(pad, np.float64), GOLD[f'lb_{tag}_pad'])


@pytest.mark.parametrize('case', SCALE_CASES, ids=[c[0] for c in SCALE_CASES])
def test_scale_coords_oracle_matches_reference(case):
    tag, s1, s0, n, seed, use_rp = case
    d = synth_boxes(n, s1, seed)
    rp = None
    if use_rp:
        _, _, ratio, pad = R.letterbox_geometry(s0, s1, auto=False)
        rp = (ratio, pad)
    out = R.scale_coords_ref(s1, d, s0, rp)
    assert np.array_equal(out, GOLD[f'sc_{tag}'])
    assert np.array_equal(R.scale_coords_ref(s1, d, s0, rp, do_round=True)[:, :4], GOLD[f'sc_{tag}_round'])


def test_resize_oracle_matches_installed_cv2():
    """Wider sweep against the binary itself where it is installed (same image here and on the GPU box)."""
    cv2 = pytest.importorskip('cv2')
    rng = np.random.default_rng(5)
    for _ in range(40):
        sh, sw = int(rng.integers(1, 300)), int(rng.integers(1, 300))
        dh, dw = int(rng.integers(1, 300)), int(rng.integers(1, 300))
        img = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
        want = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(R.resize_linear_u8(img, dw, dh), want), (sh, sw, dh, dw)
    img = rng.integers(0, 256, (200, 120, 3), dtype=np.uint8)    # exact 2x: box average
    assert np.array_equal(R.resize_linear_u8(img, 60, 100), cv2.resize(img, (60, 100), interpolation=cv2.INTER_LINEAR))
