"""GPU parity of the evaluation-side kernels: yb_box_iou and yb_match_detections against the
reference's golden output (tests/golden/eval_match.npz, produced by the reference's own test.py lines) and the oracle
(oracle/eval_ref.py).  Integer / boolean work: bit-exact."""
import numpy as np
import pytest
import torch

from oracle.eval_ref import box_iou_ref, match_image_ref
from oracle.make_golden_tta import eval_case
from oracle.synth import seeded_state_dict, synth_image

pytestmark = pytest.mark.gpu
IMG1 = (384, 640)


def test_box_iou_bit_exact(gold):
    from yolo_series_b200.evaluate import box_iou
    g = gold('eval_match.npz')
    dets, counts, targets, _ = eval_case(31, img1=IMG1)
    t = targets[targets[:, 0] == 0, 2:6]
    tb = torch.cat((t[:, :2] - t[:, 2:] / 2, t[:, :2] + t[:, 2:] / 2), 1)
    got = box_iou(dets[0, :int(counts[0]), :4].cuda(), tb.cuda()).cpu().numpy()
    assert np.array_equal(got, g['box_iou'])
    gen = torch.Generator().manual_seed(5)
    a = torch.rand((777, 4), generator=gen) * 300
    a[:, 2:] += a[:, :2]
    b = torch.rand((1301, 4), generator=gen) * 300
    b[:, 2:] += b[:, :2]
    b[7] = a[3]                                     # IoU exactly 1
    got = box_iou(a.cuda(), b.cuda()).cpu().numpy()
    assert np.array_equal(got, box_iou_ref(a.numpy(), b.numpy()))
    assert got[3, 7] == 1.0
    assert box_iou(a[:0].cuda(), b.cuda()).shape == (0, 1301)


def test_match_batch_bit_exact_vs_reference(gold):
    from yolo_series_b200.evaluate import check_match_status, match_batch
    g = gold('eval_match.npz')
    dets, counts, targets, metas = eval_case(31, img1=IMG1)
    iouv = torch.linspace(0.5, 0.95, 10)
    before = dets.clone()
    d = dets.cuda()
    correct, detected = match_batch(d, counts.cuda(), targets, IMG1, metas, iouv, return_detected=True)
    check_match_status(correct)
    assert np.array_equal(correct.cpu().numpy().astype(np.uint8), g['correct'])
    assert torch.equal(d.cpu(), before)             # predictions are not modified (test.py clones into predn)
    assert int(detected.sum()) == int(correct[..., 0].sum())


@pytest.mark.parametrize('seed,B,max_det,nc', [(1, 8, 300, 3), (2, 32, 300, 80), (3, 5, 1000, 1)])
def test_match_batch_vs_oracle(seed, B, max_det, nc):
    from yolo_series_b200.evaluate import check_match_status, match_batch
    dets, counts, targets, metas = eval_case(seed, B=B, max_det=max_det, nc=nc, img1=IMG1)
    iouv = torch.linspace(0.5, 0.95, 10)
    correct, detected = match_batch(dets.cuda(), counts.cuda(), targets, IMG1, metas, iouv, return_detected=True)
    check_match_status(correct)
    c, dt = correct.cpu().numpy(), detected.cpu().numpy()
    off = 0
    for b in range(B):
        n = int(counts[b])
        labels = targets[targets[:, 0] == b, 1:].numpy()
        m = metas[b]
        rc, rd = match_image_ref(dets[b, :n].numpy(), labels, m['ratio'][0], m['pad'], m['shape0'], iouv.numpy())
        assert np.array_equal(c[b, :n], rc), b
        assert not c[b, n:].any()
        assert np.array_equal(dt[off:off + labels.shape[0]], rd), b
        off += labels.shape[0]


def test_match_batch_rejects_ungrouped_targets():
    from yolo_series_b200.evaluate import check_match_status, match_batch
    dets, counts, targets, metas = eval_case(31, img1=IMG1)
    bad = targets.flip(0).contiguous()
    correct = match_batch(dets.cuda(), counts.cuda(), bad, IMG1, metas, torch.linspace(0.5, 0.95, 10))
    with pytest.raises(ValueError):
        check_match_status(correct)


def test_match_batch_no_targets():
    from yolo_series_b200.evaluate import match_batch
    dets, counts, _, metas = eval_case(31, img1=IMG1)
    correct = match_batch(dets.cuda(), counts.cuda(), torch.zeros((0, 6)), IMG1, metas, torch.linspace(0.5, 0.95, 10))
    assert not correct.any()


def test_autoshape_equals_manual_pipeline():
    """autoShape (common.py:865-932) = letterbox(auto=False) -> model -> NMS -> scale_coords, frames of mixed sizes."""
    from yolo_series_b200 import Model, non_max_suppression
    from yolo_series_b200.evaluate import NMS, autoShape
    from yolo_series_b200.preprocess import letterbox, scale_coords
    m = Model('yolov7-tiny')
    m.load_state_dict(seeded_state_dict(m, seed=0))
    m = m.set_precision('fp32').cuda().eval()
    gen = torch.Generator().manual_seed(9)
    frames = [torch.randint(0, 256, s + (3,), generator=gen, dtype=torch.uint8).numpy() for s in ((200, 300), (240, 180))]
    res = autoShape(m)(frames, size=160)
    assert res.n == 2 and len(res.xywhn) == 2
    shape1 = res.s[2:]
    for i, f in enumerate(frames):
        x = letterbox(f, new_shape=tuple(shape1), auto=False)[0].permute(2, 0, 1)[None].contiguous()
        pred = m(x)[0]
        det = non_max_suppression(pred, 0.25, 0.45)[0].clone()
        scale_coords(shape1, det[:, :4], f.shape[:2])
        assert torch.equal(res.pred[i], det), i
        nm = NMS()
        assert torch.equal(nm((pred, None))[0], non_max_suppression(pred, 0.25, 0.45)[0])
