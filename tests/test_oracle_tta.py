"""Pins the TTA / ensemble oracle (oracle/tta_ref.py) against the real reference's outputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle.nms_ref import non_max_suppression_ref
from oracle.synth import seeded_state_dict, synth_image
from oracle.tta_ref import ensemble_forward_ref, scale_img_ref, tta_forward_ref
from yolo_series_b200 import Model, zoo


def test_scale_img_matches_reference_golden(gold):
    g = gold('tta_scale_img.npz')
    xs = synth_image(2, 96, 136, seed=7)
    for ratio in (0.83, 0.67):
        for fi in (None, 2, 3):
            got = scale_img_ref(xs, ratio, gs=32, flip=fi).numpy()
            assert np.array_equal(got, g[f'r{ratio}_f{fi}']), (ratio, fi)
    # an exact 2x reduction goes through a different ATen loop specialisation: same formula, 1-ulp contraction differences
    assert np.abs(scale_img_ref(xs, 0.5, same_shape=True).numpy() - g['r0.5_same']).max() <= 2.4e-7


@pytest.mark.parametrize('shape,ratio', [((1, 3, 64, 64), 0.83), ((2, 1, 37, 53), 0.67), ((1, 2, 640, 640), 0.83),
                                         ((1, 1, 33, 1000), 0.67), ((1, 1, 128, 160), 1.7)])
def test_scale_img_is_atens_bilinear(shape, ratio):
    """Against the torch build in this image on shapes the goldens do not cover.  ATen picks its loop specialisation
    (and with it the compiler's FMA contraction) by size and layout, so small / single-channel inputs agree to 2 ulp
    of the [0, 1] value range rather than bit for bit; the TTA shapes themselves are bit-exact (goldens above)."""
    x = torch.rand(shape, generator=torch.Generator().manual_seed(3))
    h, w = shape[2:]
    s = (int(h * ratio), int(w * ratio))
    ref = F.interpolate(x, size=s, mode='bilinear', align_corners=False)
    got = scale_img_ref(x, ratio, same_shape=False, gs=1)
    assert got.shape[2] >= s[0] and got.shape[3] >= s[1]
    assert (got[:, :, :s[0], :s[1]] - ref).abs().max() <= 2.4e-7


def test_tta_forward_matches_reference(gold):
    g = gold('tta_tiny.npz')
    cfg = zoo.get_cfg('yolov7-tiny')
    sd = seeded_state_dict(Model(cfg), seed=0)
    pred = tta_forward_ref(cfg, sd, synth_image(2, 128, 160, seed=2))
    ref = torch.from_numpy(g['pred'])
    assert pred.shape == ref.shape
    assert torch.allclose(pred, ref, rtol=2e-5, atol=3e-4), (pred - ref).abs().max()
    dets = [non_max_suppression_ref(ref[i:i + 1], 0.25, 0.45)[0] for i in range(ref.shape[0])]
    assert [d.shape[0] for d in dets] == g['counts'].tolist()
    assert np.array_equal(torch.cat(dets, 0).numpy(), g['dets'])


def test_ensemble_forward_matches_reference(gold):
    g = gold('tta_ensemble.npz')
    cfg = zoo.get_cfg('yolov7-tiny')
    members = [(cfg, seeded_state_dict(Model(cfg), seed=s)) for s in (0, 5)]
    pred = ensemble_forward_ref(members, synth_image(2, 128, 160, seed=2))
    ref = torch.from_numpy(g['pred'])
    assert pred.shape == ref.shape
    assert torch.allclose(pred, ref, rtol=2e-5, atol=3e-4), (pred - ref).abs().max()
