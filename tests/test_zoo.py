"""zoo.py generates the same graphs as the reference YAML files."""
import pytest
from conftest import HAVE_REF

from yolo_series_b200 import zoo


def test_zoo_shapes():
    n = {k: len(zoo.get_cfg(k)['backbone']) + len(zoo.get_cfg(k)['head']) for k in zoo.ZOO}
    assert n == {'yolov7': 106, 'yolov7-tiny': 78, 'yolov7-w6': 123, 'yolov7-w6-custom': 123, 'yolov7-e6e': 266}
    assert zoo.get_cfg('yolov7-w6-custom')['nc'] == 83


@pytest.mark.skipif(not HAVE_REF, reason='/root/reference not mounted')
@pytest.mark.parametrize('name', sorted(zoo.ZOO))
def test_zoo_equals_reference_yaml(name):
    import yaml
    with open(f'/root/reference/cfg/training/{name}.yaml') as f:
        ref = yaml.safe_load(f)
    mine = zoo.get_cfg(name)
    for k in ('nc', 'depth_multiple', 'width_multiple', 'anchors', 'backbone', 'head'):
        assert ref[k] == mine[k], k


@pytest.mark.skipif(not HAVE_REF, reason='/root/reference not mounted')
@pytest.mark.parametrize('name', ['yolov7-tiny', 'yolov7', 'yolov7-w6-custom'])
def test_accelerate_lowers_a_real_reference_model(name):
    """``accelerate(ref_model)`` duck-types the reference's own module classes: lowering a reference
    ``Model`` -- as built, and after its own ``fuse()`` (Conv+BN folded, RepConv re-parameterised, Implicit folded,
    forward swapped for fuseforward) -- must give the same launch list and the same folded weights as lowering this
    package's Model with the same state_dict."""
    import torch

    from oracle.make_golden import import_reference
    from oracle.synth import seeded_state_dict
    from yolo_series_b200 import Model, accelerate, plan, zoo
    RefModel, _ = import_reference()
    cfg = zoo.get_cfg(name)
    ours = Model(cfg)
    sd = seeded_state_dict(ours, seed=0)
    ours.load_state_dict(sd)
    with torch.no_grad():
        want = plan.lower(ours.model, in_ch=3, fused=False)
        for fuse in (False, True):
            ref = RefModel(cfg).eval()
            ref.load_state_dict(sd)
            if fuse:
                ref.fuse()
            acc = accelerate(ref)
            assert acc._is_ref_fused() == fuse and acc.stride.tolist() == ours.stride.tolist()
            assert next(acc.parameters()) is not None          # detect.py:73
            got = plan.lower(acc._layers(), in_ch=3, fused=acc._is_ref_fused())
            assert len(got.ops) == len(want.ops)
            for a, b in zip(want.ops, got.ops):
                assert (a.kind, a.ins, a.outs, a.upsample) == (b.kind, b.ins, b.outs, b.upsample), (a.name, b.name)
                if a.kind == 1:
                    assert (a.conv.k, a.conv.s, a.conv.act, a.conv.slope) == (b.conv.k, b.conv.s, b.conv.act, b.conv.slope)
                    assert torch.allclose(a.conv.weight, b.conv.weight, rtol=1e-5, atol=1e-6), a.name
                    assert torch.allclose(a.conv.bias, b.conv.bias, rtol=1e-5, atol=1e-6), a.name
