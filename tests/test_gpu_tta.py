"""GPU parity of the TTA / ensemble callers: yb_scale_img, yb_decode_tta, Model.forward(augment=True),
Ensemble.forward -- against the reference's golden outputs (tests/golden/tta_*.npz) and the oracle (oracle/tta_ref.py).
Tolerances: scale_img bit-exact; fp32 mode boxes <= 1e-2 px / scores <= 1e-3 (north_star); the de-scale / de-flip of the
decode bit-exact against the same arithmetic applied to yb_decode's output."""
import numpy as np
import pytest
import torch

from oracle.nms_ref import non_max_suppression_ref
from oracle.synth import seeded_state_dict, synth_image
from oracle.tta_ref import scale_img_ref

pytestmark = pytest.mark.gpu


def _build(name, mode, seed=0):
    from yolo_series_b200 import Model
    m = Model(name)
    m.load_state_dict(seeded_state_dict(m, seed=seed))
    m.set_precision(mode)
    return m.cuda().eval()


def test_scale_img_bit_exact_vs_reference_golden(gold):
    g = gold('tta_scale_img.npz')
    m = _build('yolov7-tiny', 'fp32')
    xs = synth_image(2, 96, 136, seed=7).cuda()
    for ratio in (0.83, 0.67):
        for fi in (None, 2, 3):
            ref = g[f'r{ratio}_f{fi}']
            h, w = int(96 * ratio), int(136 * ratio)
            got = m._scaled_input(xs, ratio, fi, (h, w), ref.shape[2:]).cpu().numpy()
            assert np.array_equal(got, ref), (ratio, fi, np.abs(got - ref).max())


@pytest.mark.parametrize('shape,ratio,flip', [((2, 3, 640, 640), 0.83, 3), ((1, 3, 640, 640), 0.67, None),
                                              ((3, 3, 1280, 1280), 0.83, 3), ((1, 12, 37, 53), 1.7, 2)])
def test_scale_img_matches_oracle(shape, ratio, flip):
    m = _build('yolov7-tiny', 'fp32')
    x = torch.rand(shape, generator=torch.Generator().manual_seed(11))
    ref = scale_img_ref(x, ratio, gs=32, flip=flip)
    h, w = int(shape[2] * ratio), int(shape[3] * ratio)
    got = m._scaled_input(x.cuda(), ratio, flip, (h, w), tuple(ref.shape[2:])).cpu()
    assert torch.equal(got, ref)


def test_scale_img_uint8_source():
    m = _build('yolov7-tiny', 'fp32')
    x = synth_image(2, 128, 160, seed=4, dtype=torch.uint8)
    ref = scale_img_ref(x.float() / 255.0, 0.83, gs=32, flip=3)
    got = m._scaled_input(x.cuda(), 0.83, 3, (106, 132), (128, 160)).cpu()
    assert torch.equal(got, ref)


def test_tta_forward_fp32_matches_reference(gold):
    from yolo_series_b200 import non_max_suppression
    g = gold('tta_tiny.npz')
    m = _build('yolov7-tiny', 'fp32')
    x = synth_image(2, 128, 160, seed=2).cuda()
    pred, raw = m(x, augment=True)
    assert raw is None
    ref = torch.from_numpy(g['pred'])
    p = pred.cpu()
    assert p.shape == ref.shape
    assert float((p[..., :4] - ref[..., :4]).abs().max()) <= 1e-2
    assert float((p[..., 4:] - ref[..., 4:]).abs().max()) <= 1e-3
    dets = non_max_suppression(pred, 0.25, 0.45)
    r = non_max_suppression_ref(p, 0.25, 0.45)
    for a, b in zip(dets, r):
        assert np.array_equal(a.cpu().numpy(), b.numpy())     # NMS bit-exact on the 3x candidate set
    assert [d.shape[0] for d in dets] == g['counts'].tolist()
    fused = m.detect(x, 0.25, 0.45, augment=True)
    for a, b in zip(fused, dets):
        assert torch.equal(a, b)


def test_tta_passes_are_descaled_decodes():
    """Pass 0 of the TTA output is the plain forward; passes 1-2 equal the plain forward of the scaled input with
    the reference's de-scale / de-flip arithmetic applied (bit for bit)."""
    m = _build('yolov7-tiny', 'fp32')
    x = synth_image(1, 128, 160, seed=2).cuda()
    pred = m(x, augment=True)[0]
    p0 = m(x)[0]
    A0 = p0.shape[1]
    assert torch.equal(pred[:, :A0], p0)
    x1 = m._scaled_input(x, 0.83, 3, (106, 132), (128, 160))
    pred = pred.cpu()
    p1 = m(x1)[0].cpu()          # the arithmetic below on the CPU: torch's CUDA `/= scalar` multiplies by a reciprocal
    p1[..., :4] /= 0.83
    p1[..., 0] = 160 - p1[..., 0]
    assert torch.equal(pred[:, A0:A0 + p1.shape[1]], p1)
    x2 = m._scaled_input(x, 0.67, None, (85, 107), (96, 128))
    p2 = m(x2)[0].cpu()
    p2[..., :4] /= 0.67
    assert torch.equal(pred[:, A0 + p1.shape[1]:], p2)


def test_tta_bf16_runs_close():
    m32, m16 = _build('yolov7-tiny', 'fp32'), _build('yolov7-tiny', 'bf16')
    x = synth_image(2, 128, 160, seed=2, dtype=torch.uint8).cuda()
    a, b = m32(x, augment=True)[0], m16(x, augment=True)[0]
    assert a.shape == b.shape
    assert float((a[..., 4:] - b[..., 4:]).abs().quantile(0.99)) <= 0.03


def test_ensemble_fp32_matches_reference(gold):
    from yolo_series_b200 import Ensemble, non_max_suppression
    g = gold('tta_ensemble.npz')
    ens = Ensemble([_build('yolov7-tiny', 'fp32', seed=s) for s in (0, 5)])
    x = synth_image(2, 128, 160, seed=2).cuda()
    pred, raw = ens(x)
    assert raw is None
    ref = torch.from_numpy(g['pred'])
    p = pred.cpu()
    assert p.shape == ref.shape
    assert float((p[..., :4] - ref[..., :4]).abs().max()) <= 1e-2
    assert float((p[..., 4:] - ref[..., 4:]).abs().max()) <= 1e-3
    for i, mdl in enumerate(ens):
        assert torch.equal(pred[:, i * 1260:(i + 1) * 1260], mdl(x)[0])
    dets = non_max_suppression(pred, 0.25, 0.45)
    assert [d.shape[0] for d in dets] == g['counts'].tolist()
    pa = ens(x, augment=True)[0]
    assert pa.shape[1] == 2 * 3276
