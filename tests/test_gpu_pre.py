"""GPU parity of the pre-/post-processing kernels through the C ABI
(yb_letterbox / yb_scale_coords): bit-exact against fixtures written by the REAL reference
(utils/datasets.py letterbox + cv2, utils/general.py scale_coords) and against the oracle on wider sweeps."""
import numpy as np
import pytest
import torch

from oracle import preprocess_ref as R
from oracle.make_golden_pre import LETTERBOX_CASES, SCALE_CASES, synth_boxes, synth_frame
from test_oracle_pre import GOLD, check_letterbox_against_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('case', LETTERBOX_CASES, ids=[c[0] for c in LETTERBOX_CASES])
def test_letterbox_matches_reference_golden(case):
    from yolo_series_b200.preprocess import letterbox
    tag, h, w, seed, kw = case
    img, ratio, pad = letterbox(synth_frame(h, w, seed), **kw)
    assert img.is_cuda and img.dtype == torch.uint8
    check_letterbox_against_golden(tag, img.cpu().numpy())
    assert np.array_equal(np.array(ratio, np.float64), GOLD[f'lb_{tag}_ratio'])
    assert np.array_equal(np.array(pad, np.float64), GOLD[f'lb_{tag}_pad'])


def test_letterbox_random_sizes_match_oracle():
    """Ragged sizes, 1-pixel sources, widths that are not a multiple of 4 (scalar store path)."""
    from yolo_series_b200.preprocess import letterbox
    rng = np.random.default_rng(11)
    for i in range(30):
        sh, sw = int(rng.integers(1, 400)), int(rng.integers(1, 400))
        new = (int(rng.integers(8, 300)), int(rng.integers(8, 300)))
        kw = dict(new_shape=new, auto=bool(i & 1), scaleup=bool(i & 2), stride=int(rng.choice([8, 32, 64])),
                  color=tuple(int(v) for v in rng.integers(0, 256, 3)))
        (nw, nh), _, _, _ = R.letterbox_geometry((sh, sw), new, kw['auto'], False, kw['scaleup'], kw['stride'])
        if nw < 1 or nh < 1:        # cv2.resize rejects an empty destination; so does yb_letterbox
            continue
        frame = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
        want, ratio, pad = R.letterbox_ref(frame, **kw)
        got, ratio2, pad2 = letterbox(frame, **kw)
        assert np.array_equal(got.cpu().numpy(), want), (sh, sw, kw)
        assert ratio2 == ratio and pad2 == pad


def test_letterbox_batch_chw_rgb_mixed_sizes():
    """One launch, frames of different sizes -> (B, 3, H, W) RGB planar = datasets.py:198-199 after letterbox."""
    from yolo_series_b200.preprocess import letterbox_batch
    sizes = [(480, 640), (720, 1280), (333, 500), (1080, 1920), (640, 640), (1280, 1280), (37, 53)]
    frames = [synth_frame(h, w, 40 + i) for i, (h, w) in enumerate(sizes)]
    x, metas = letterbox_batch(frames, new_shape=640, auto=False)
    assert tuple(x.shape) == (len(sizes), 3, 640, 640) and x.dtype == torch.uint8
    xs = x.cpu().numpy()
    for i, f in enumerate(frames):
        want, ratio, pad = R.letterbox_ref(f, 640, auto=False)
        assert np.array_equal(xs[i], R.to_model_input(want)), sizes[i]
        assert metas[i]['ratio'] == ratio and metas[i]['pad'] == pad and metas[i]['shape0'] == sizes[i]


def test_letterbox_batch_more_than_one_launch_group():
    """B > 32 splits into several launches; pitched (non-contiguous rows) device sources."""
    from yolo_series_b200.preprocess import letterbox_batch
    rng = np.random.default_rng(3)
    big = torch.from_numpy(rng.integers(0, 256, (40, 100, 160, 3), dtype=np.uint8)).cuda()
    frames = [big[i, :90, :120] for i in range(40)]         # row pitch 480 B, width 120
    x, _ = letterbox_batch(frames, new_shape=64, auto=False, swap_rb=False)
    xs = x.cpu().numpy()
    for i in (0, 17, 31, 32, 39):
        want, _, _ = R.letterbox_ref(frames[i].cpu().numpy(), 64, auto=False)
        assert np.array_equal(xs[i], want.transpose(2, 0, 1))


def test_letterbox_errors():
    from yolo_series_b200.preprocess import letterbox, letterbox_batch
    with pytest.raises(ValueError):
        letterbox(np.zeros((4, 4), np.uint8))
    with pytest.raises(ValueError):
        letterbox(np.zeros((4, 4, 3), np.float32))
    with pytest.raises(ValueError):
        letterbox_batch([np.zeros((40, 80, 3), np.uint8), np.zeros((80, 40, 3), np.uint8)], 64, auto=True)


@pytest.mark.parametrize('case', SCALE_CASES, ids=[c[0] for c in SCALE_CASES])
def test_scale_coords_matches_reference_golden(case):
    from yolo_series_b200.preprocess import scale_coords
    tag, s1, s0, n, seed, use_rp = case
    d = torch.from_numpy(synth_boxes(n, s1, seed)).cuda()
    rp = None
    if use_rp:
        _, _, ratio, pad = R.letterbox_geometry(s0, s1, auto=False)
        rp = (ratio, pad)
    ret = scale_coords(s1, d[:, :4], s0, rp)           # a strided view, edited in place like detect.py:183
    assert ret.data_ptr() == d.data_ptr()
    assert np.array_equal(d.cpu().numpy(), GOLD[f'sc_{tag}'])
    assert np.array_equal(d[:, :4].round().cpu().numpy(), GOLD[f'sc_{tag}_round'])


def test_clip_coords_and_empty():
    from yolo_series_b200.preprocess import clip_coords, scale_coords
    b = torch.tensor([[-5., -1., 700., 500.], [10., 20., 30., 40.], [float('nan'), 1., 2., 3.]]).cuda()
    clip_coords(b, (480, 640))
    got = b.cpu()
    assert got[0].tolist() == [0., 0., 640., 480.] and got[1].tolist() == [10., 20., 30., 40.]
    assert torch.isnan(got[2, 0]) and got[2, 1:].tolist() == [1., 2., 3.]
    e = torch.zeros((0, 6)).cuda()
    assert scale_coords((640, 640), e, (480, 640)).shape == (0, 6)


def test_rescale_detections_packed_counts():
    """All images of the packed NMS output in one launch; rows past counts[b] are left untouched."""
    from yolo_series_b200.preprocess import rescale_detections
    s1 = (640, 640)
    shapes = [(480, 640), (720, 1280), (333, 500), (1080, 1920)]
    B, max_det = len(shapes), 300
    host = np.stack([synth_boxes(max_det, s1, 60 + b) for b in range(B)])
    counts = np.array([300, 0, 17, 299], np.int32)
    metas = []
    for s0 in shapes:
        _, _, ratio, pad = R.letterbox_geometry(s0, s1, auto=False)
        metas.append(dict(ratio=ratio, pad=pad, shape0=s0))
    for use_rp in (False, True):
        out = torch.from_numpy(host.copy()).cuda()
        rescale_detections(out, torch.from_numpy(counts).cuda(), s1, metas, round_=True, use_ratio_pad=use_rp)
        got = out.cpu().numpy()
        for b, s0 in enumerate(shapes):
            rp = (metas[b]['ratio'], metas[b]['pad']) if use_rp else None
            want = R.scale_coords_ref(s1, host[b], s0, rp, do_round=True)
            n = counts[b]
            assert np.array_equal(got[b, :n], want[:n])
            assert np.array_equal(got[b, n:], host[b, n:])
