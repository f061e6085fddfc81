"""Pins the oracle forward (oracle/ref_forward.py) against the real reference's outputs."""
import numpy as np
import pytest
import torch

from oracle.make_golden import FULL_CASES, FWD_CASES
from oracle.nms_ref import non_max_suppression_ref
from oracle.ref_forward import ref_forward
from oracle.synth import seeded_state_dict, synth_image
from yolo_series_b200 import Model, zoo


@pytest.mark.parametrize('case', FWD_CASES, ids=[c[0] for c in FWD_CASES])
def test_forward_matches_reference(case, gold):
    name, B, H, W, conf, iou = case
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    cfg = zoo.get_cfg(name)
    sd = seeded_state_dict(Model(cfg), seed=0)
    pred, raws = ref_forward(cfg, sd, synth_image(B, H, W, seed=2), fused=True)
    assert list(pred.shape) == g['shape'].tolist()
    ref = torch.from_numpy(g['pred'])
    # fold association differs from the reference's mm-based fold only at rounding level
    assert torch.allclose(pred, ref, rtol=2e-5, atol=2e-4), (pred - ref).abs().max()
    assert torch.allclose(raws[0], torch.from_numpy(g['raw0']), rtol=1e-4, atol=1e-4)
    # NMS on the *reference's* pred must reproduce the reference's detections bit for bit
    dets = non_max_suppression_ref(ref, float(g['conf']), float(g['iou']))
    assert [d.shape[0] for d in dets] == g['counts'].tolist()
    assert np.array_equal(torch.cat(dets, 0).numpy(), g['dets'])


@pytest.mark.parametrize('case', FULL_CASES, ids=[f'{c[0]}_{c[2]}' for c in FULL_CASES])
def test_forward_matches_reference_at_full_size(case, gold):
    """BASELINE.json configs[1]-[3] at 640 / 1280: the oracle forward against the reference's own output (every 37th
    row of pred is stored).  The fold-association rounding grows with the coordinate magnitude (1280-px boxes)."""
    name, B, H, W, conf, iou = case
    g = gold(f'fwd_{name}_{H}x{W}.npz')
    cfg = zoo.get_cfg(name)
    sd = seeded_state_dict(Model(cfg), seed=0)
    pred, _ = ref_forward(cfg, sd, synth_image(B, H, W, seed=2), fused=True)
    assert list(pred.shape) == g['shape'].tolist()
    sub, ref = pred[:, ::37], torch.from_numpy(g['pred_sub'])
    assert float((sub[..., :4] - ref[..., :4]).abs().max()) <= 5e-3
    assert float((sub[..., 4:] - ref[..., 4:]).abs().max()) <= 1e-5
    assert int((pred[..., 4] > conf).sum()) == int(g['cand'][0])
