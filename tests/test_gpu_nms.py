"""GPU parity: CUDA NMS / decode (through the C ABI) vs the reference's golden outputs and the oracle.
Bar: bit-exact rows (values, order, counts)."""
import numpy as np
import pytest
import torch

from oracle.make_golden import NMS_CASES, WIDE_CASES
from oracle.nms_ref import non_max_suppression_ref
from oracle.synth import synth_pred, synth_pred_wide

pytestmark = pytest.mark.gpu


def _flat(dets):
    return torch.cat([d.reshape(-1, 6).float().cpu() for d in dets], 0).numpy()


@pytest.mark.parametrize('case', NMS_CASES, ids=[c[0] for c in NMS_CASES])
def test_nms_equals_reference_golden(case, gold):
    from yolo_series_b200 import non_max_suppression
    tag, B, A, nc, seed, p_obj, kw = case
    g = gold(f'nms_{tag}.npz')
    pred = synth_pred(B, A, nc, seed=seed, p_obj=p_obj).cuda()
    out = non_max_suppression(pred, **kw)
    assert [o.shape[0] for o in out] == g['counts'].tolist()
    assert np.array_equal(_flat(out), g['dets'])


@pytest.mark.parametrize('case', WIDE_CASES, ids=[c[0] for c in WIDE_CASES])
def test_nms_cross_class_suppression(case, gold):
    """Boxes of 6 000-9 000 px: after the 4096-px class offset (general.py:702-704) neighbouring classes overlap and
    suppress each other.  k_sort_nms chains the kept set per class bucket only while the un-offset x extent of
    everything seen stays < 4000 px and must drop that mode here -- at once ('wide') or after several chunks of narrow
    boxes ('wide_mixed').  Bit-exact against the real reference's golden and the oracle; tests/test_oracle_nms.py proves
    from the oracle side that suppressions do cross classes in these inputs."""
    from yolo_series_b200 import non_max_suppression
    tag, B, A, nc, seed, mixed, kw = case
    g = gold(f'nms_{tag}.npz')
    pred = synth_pred_wide(B, A, nc, seed=seed, mixed=mixed)
    out = non_max_suppression(pred.cuda(), **kw)
    assert [o.shape[0] for o in out] == g['counts'].tolist()
    assert np.array_equal(_flat(out), g['dets'])
    for extra in (dict(max_det=1000), dict(max_det=20), dict(agnostic=True), dict(multi_label=True)):
        k2 = {**kw, **extra}
        ref = non_max_suppression_ref(pred, **k2)
        got = non_max_suppression(pred.cuda(), **k2)
        assert [o.shape[0] for o in got] == [r.shape[0] for r in ref], extra
        assert np.array_equal(_flat(got), _flat(ref)), extra


@pytest.mark.parametrize('kw', [
    dict(conf_thres=0.001, iou_thres=0.65, max_det=1000),
    dict(conf_thres=0.001, iou_thres=0.65, multi_label=True, max_det=300),
    dict(conf_thres=0.25, iou_thres=0.45, agnostic=True, max_det=50),
    dict(conf_thres=0.1, iou_thres=0.3, classes=[1, 2, 3, 79]),
    dict(conf_thres=0.001, iou_thres=0.65, max_det=1000, max_nms=2000),
], ids=['maxdet1000', 'multilabel', 'agnostic50', 'classes', 'maxnms2000'])
def test_nms_equals_oracle_dense(kw):
    from yolo_series_b200 import non_max_suppression
    pred = synth_pred(3, 20000, 80, seed=101, p_obj=0.3, clusters=(100, 200), extent=1280.0)
    ref = non_max_suppression_ref(pred, **kw)
    out = non_max_suppression(pred.cuda(), **kw)
    assert [o.shape[0] for o in out] == [r.shape[0] for r in ref]
    assert np.array_equal(_flat(out), _flat(ref))


def test_nms_stress_density_100k():
    """BASELINE.json config 5 density (100k rows/image, 80 classes, max_det 1000) on 2 images."""
    from yolo_series_b200 import non_max_suppression
    pred = synth_pred(2, 100000, 80, seed=7, p_obj=0.3, clusters=(400, 800), extent=1280.0)
    kw = dict(conf_thres=0.001, iou_thres=0.65, max_det=1000)
    ref = non_max_suppression_ref(pred, **kw)
    out = non_max_suppression(pred.cuda(), **kw)
    assert [o.shape[0] for o in out] == [r.shape[0] for r in ref]
    assert np.array_equal(_flat(out), _flat(ref))


def test_nms_suppression_heavy_never_reaches_max_det():
    """All boxes of one class in a few tight clusters: the kept set stays tiny and every candidate must be scanned."""
    from yolo_series_b200 import non_max_suppression
    pred = synth_pred(2, 30000, 1, seed=9, p_obj=1.0, clusters=(5, 8), extent=640.0, jitter=2.0)
    kw = dict(conf_thres=0.05, iou_thres=0.45)
    ref = non_max_suppression_ref(pred, **kw)
    out = non_max_suppression(pred.cuda(), **kw)
    assert np.array_equal(_flat(out), _flat(ref))


@pytest.mark.parametrize('mode', ['all_equal', 'few_levels', 'two_level_refine'])
def test_nms_score_ties_across_bands(mode):
    """The sort+NMS kernel takes candidates in score bands picked by a radix select; equal scores must still come
    out in index order, whether the tie fits one band, spans the select's refinement levels or overflows a band
    (more than 4096 equal scores -> full sort fallback)."""
    from yolo_series_b200 import non_max_suppression
    pred = synth_pred(2, 12000, 3, seed=21, p_obj=1.0, extent=2000.0)
    g = torch.Generator().manual_seed(5)
    if mode == 'all_equal':            # 12000 identical scores: one tie larger than a band
        pred[..., 4] = 1.0
        pred[..., 5:] = 0.5
    elif mode == 'few_levels':         # 4 distinct scores, 3000 boxes each
        pred[..., 4] = 1.0
        lv = torch.tensor([0.9, 0.7, 0.5, 0.3])[torch.randint(0, 4, (2, 12000), generator=g)]
        pred[..., 5:] = lv[..., None]
    else:                              # scores differing only in the low mantissa bits
        pred[..., 4] = 1.0
        base = torch.tensor(0.75).view(torch.int32)
        bits = base + torch.randint(0, 200, (2, 12000), generator=g, dtype=torch.int32)
        pred[..., 5:] = bits.view(torch.float32)[..., None]
    for kw in (dict(conf_thres=0.001, iou_thres=0.65, max_det=300), dict(conf_thres=0.001, iou_thres=0.2, max_det=3000)):
        ref = non_max_suppression_ref(pred, **kw)
        out = non_max_suppression(pred.cuda(), **kw)
        assert [o.shape[0] for o in out] == [r.shape[0] for r in ref]
        assert np.array_equal(_flat(out), _flat(ref))


def test_nms_edge_cases():
    from yolo_series_b200 import non_max_suppression
    z = torch.zeros((2, 64, 85)).cuda()
    out = non_max_suppression(z)
    assert [tuple(o.shape) for o in out] == [(0, 6), (0, 6)]
    out[0] = out[0] + 1  # independent tensors
    # ragged: image 0 empty, image 1 dense; labels appended (general.py:656-662)
    pred = synth_pred(2, 500, 80, seed=3, p_obj=0.5)
    pred[0, :, 4] = 0
    labels = [torch.tensor([[3., 100., 100., 50., 80.]]), torch.zeros((0, 5))]
    ref = non_max_suppression_ref(pred, 0.25, 0.45, labels=labels)
    got = non_max_suppression(pred.cuda(), 0.25, 0.45, labels=labels)
    assert np.array_equal(_flat(got), _flat(ref))
    assert got[0].shape[0] == 1 and float(got[0][0, 5]) == 3.0


def test_nms_non_finite_boxes_follow_the_reference():
    """Rows whose box coordinates are NaN / inf but whose scores pass the filter (a NaN box logit next to healthy
    obj / class logits): torchvision's kernel uses std::max / std::min (first operand on an unordered compare), the CUDA
    kernel IEEE maxNum / minNum -- the intermediate overlaps differ, but a NaN coordinate makes that box's area and
    hence every IoU it takes part in NaN, `NaN > thr` is false either way: such a row neither suppresses nor is
    suppressed.  The kept rows must equal the oracle's, NaN for NaN."""
    from yolo_series_b200 import non_max_suppression
    pred = synth_pred(3, 3000, 80, seed=17, p_obj=0.4, clusters=(20, 40))
    g = torch.Generator().manual_seed(1)
    for b in range(3):
        idx = torch.randperm(3000, generator=g)[:60]
        pred[b, idx[:15], 0] = float('nan')          # x centre
        pred[b, idx[15:30], 3] = float('nan')        # height
        pred[b, idx[30:40], 2] = float('inf')        # width: x1 = -inf, x2 = +inf, area inf, IoU = NaN or 0
        pred[b, idx[40:50], 1] = float('-inf')       # y centre
        pred[b, idx[50:60], :4] = float('nan')
        pred[b, idx, 4] = 0.9                        # they all pass the filter and rank high
    for kw in (dict(conf_thres=0.25, iou_thres=0.45), dict(conf_thres=0.001, iou_thres=0.65, multi_label=True),
               dict(conf_thres=0.25, iou_thres=0.45, agnostic=True)):
        ref = non_max_suppression_ref(pred, **kw)
        got = non_max_suppression(pred.cuda(), **kw)
        assert [r.shape[0] for r in ref] == [t.shape[0] for t in got], kw
        a, r = _flat(got), _flat(ref)
        assert np.array_equal(a, r, equal_nan=True), kw      # NaN payload bits differ between the CPU and the GPU
        assert np.array_equal(np.isnan(a), np.isnan(r)) and np.array_equal(np.isinf(a), np.isinf(r))
        assert np.isnan(a).any()                      # not vacuous: non-finite rows are among the kept ones


def test_decode_matches_torch_and_fused_path():
    """decode: <= 2 ulp from the CPU sigmoid path; fused decode+NMS == decode -> NMS."""
    from yolo_series_b200 import ops
    from yolo_series_b200.nms import _run_nms
    g = torch.Generator().manual_seed(0)
    na, nc = 3, 80
    no = nc + 5
    strides = [8., 16., 32.]
    ag = [[[12, 16], [19, 36], [40, 28]], [[36, 75], [76, 55], [72, 146]], [[142, 110], [192, 243], [459, 401]]]
    B = 2
    logits = [torch.randn((B, 32 // (2 ** i), 48 // (2 ** i), 256), generator=g) * 2 - 1 for i in range(3)]
    # CPU restatement of yolo.py:153-156
    zs, raws = [], []
    for i, t in enumerate(logits):
        ny, nx = t.shape[1:3]
        x = t[..., :na * no].view(B, ny, nx, na, no).permute(0, 3, 1, 2, 4).contiguous()
        raws.append(x)
        yv, xv = torch.meshgrid(torch.arange(ny), torch.arange(nx), indexing='ij')
        grid = torch.stack((xv, yv), 2).view(1, 1, ny, nx, 2).float()
        y = x.sigmoid()
        y[..., 0:2] = (y[..., 0:2] * 2. - 0.5 + grid) * strides[i]
        y[..., 2:4] = (y[..., 2:4] * 2) ** 2 * torch.tensor(ag[i]).float().view(1, na, 1, 1, 2)
        zs.append(y.view(B, -1, no))
    ref = torch.cat(zs, 1)
    pred, raw_gpu, lv = ops.decode([t.cuda() for t in logits], strides, ag, na, no)
    for a, b in zip(raw_gpu, raws):
        assert torch.equal(a.cpu(), b)
    p = pred.cpu()
    # obj/cls are the bare sigmoid: <= 2 ulp of fp32 (values in (0,1): 1 ulp <= 6e-8)
    assert float((p[..., 4:] - ref[..., 4:]).abs().max()) <= 2 * 6e-8
    # xy/wh: the same 2-ulp sigmoid error scaled by 2*stride / 4*anchor*(2p): <= 2e-4 px here
    assert torch.allclose(p[..., :4], ref[..., :4], rtol=1e-6, atol=2e-4)
    # fused path: identical rows to NMS over the GPU-decoded pred
    from yolo_series_b200 import non_max_suppression
    a = non_max_suppression(pred, 0.05, 0.5, max_det=300)
    b = _run_nms(None, B, pred.shape[1], no, 0.05, 0.5, None, False, False, 30000, 300, pred.device, levels=lv)
    assert [x.shape[0] for x in a] == [x.shape[0] for x in b] and sum(x.shape[0] for x in a) > 0
    assert np.array_equal(_flat(a), _flat(b))
    # and NMS keep rows are bit-exact vs the oracle when fed the same (GPU-decoded) tensor
    r = non_max_suppression_ref(pred.cpu(), 0.05, 0.5)
    assert np.array_equal(_flat(a), _flat(r))


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16], ids=['f32', 'bf16'])
def test_fused_multi_label_filter_gate_is_exact(dtype):
    """multi_label=True (test.py:126, general.py:679-681) through the fused path: k_filter_cells<ML> walks one row
    per thread and skips classes below a per-row logit gate derived from conf_thres / obj.  The gate must never
    drop a candidate: fused result == stand-alone NMS over the decoded pred (evaluates every class) == oracle, for
    thresholds from 'almost everything passes' to 'almost nothing', logits placed right at the gate, saturated
    and non-finite logits, and a class whitelist."""
    from yolo_series_b200 import non_max_suppression, ops
    from yolo_series_b200.nms import _run_nms
    g = torch.Generator().manual_seed(3)
    na, nc = 3, 80
    no = nc + 5
    strides = [8., 16., 32.]
    ag = [[[12, 16], [19, 36], [40, 28]], [[36, 75], [76, 55], [72, 146]], [[142, 110], [192, 243], [459, 401]]]
    B = 2
    logits = [torch.randn((B, 32 // (2 ** i), 48 // (2 ** i), 256), generator=g) * 2.5 - 1 for i in range(3)]
    # adversarial rows: class logits exactly at / one ulp around logit(thr / obj), huge and tiny logits, NaN / inf
    t = logits[0]
    t[0, 0, 0, 5:85] = 40.0
    t[0, 0, 1, 5:85] = -40.0
    t[0, 0, 2, 4] = 40.0                                   # obj == 1
    t[0, 0, 3, 4] = float(torch.logit(torch.tensor(0.05)))  # obj right at a threshold used below
    t[0, 0, 4, 7] = float('nan')
    t[0, 0, 5, 9] = float('inf')
    t[0, 0, 6, 11] = float('-inf')
    for k, thr in enumerate((0.001, 0.05, 0.25)):           # logits at the exact gate of row (0, 1, k)
        obj = torch.sigmoid(t[0, 1, k, 4])
        q = torch.tensor(thr) / obj
        if float(q) < 1:
            gl = torch.log(q / (1 - q))
            t[0, 1, k, 5:45] = gl
            t[0, 1, k, 45:65] = torch.nextafter(gl, torch.tensor(10.0))
            t[0, 1, k, 65:85] = torch.nextafter(gl, torch.tensor(-10.0))
    dev_logits = [x.cuda().to(dtype) for x in logits]       # kept alive: lv holds raw pointers into them
    pred, _, lv = ops.decode(dev_logits, strides, ag, na, no)
    for kw in (dict(conf_thres=0.001, iou_thres=0.65), dict(conf_thres=0.05, iou_thres=0.5),
               dict(conf_thres=0.25, iou_thres=0.45), dict(conf_thres=0.9, iou_thres=0.45),
               dict(conf_thres=0.05, iou_thres=0.5, classes=[0, 3, 79])):
        a = non_max_suppression(pred, multi_label=True, max_det=300, **kw)
        b = _run_nms(None, B, pred.shape[1], no, kw['conf_thres'], kw['iou_thres'], kw.get('classes'), False, True, 30000,
                     300, pred.device, levels=lv)
        assert [x.shape[0] for x in a] == [x.shape[0] for x in b], kw
        assert np.array_equal(_flat(a), _flat(b)), kw
        r = non_max_suppression_ref(pred.cpu(), multi_label=True, **kw)
        assert np.array_equal(_flat(a), _flat(r)), kw
    assert sum(x.shape[0] for x in a) > 0
