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


OTHER_CFGS = [            # (path under /root/reference/cfg, nodes, convs after lowering (merged siblings count once), head)
    ('training/yolov7x.yaml', 'IDetect'),
    ('training/yolov7-e6.yaml', 'IAuxDetect'),
    ('training/yolov7-d6.yaml', 'IAuxDetect'),
    ('training/yolov7-custom.yaml', 'IDetect'),
    ('deploy/yolov7.yaml', 'Detect'),
    ('deploy/yolov7-tiny-silu.yaml', 'Detect'),
    ('deploy/yolov7-w6.yaml', 'Detect'),
    ('deploy/yolov7-e6e.yaml', 'Detect'),
]


@pytest.mark.skipif(not HAVE_REF, reason='/root/reference not mounted')
@pytest.mark.parametrize('rel,head', OTHER_CFGS, ids=[c[0] for c in OTHER_CFGS])
def test_other_reference_cfgs_build_lower_and_match_the_reference(rel, head):
    """The op set of SURVEY.md A.1 covers every yolov7 training and deploy cfg, not just the five zoo names:
    ``Model(path-to-yaml)`` must build with the reference's state_dict layout, lower without a single copy op, and the
    oracle forward on that cfg must equal the real reference's forward (which pins the oracle on these graphs too)."""
    import torch
    import yaml

    from oracle.make_golden import import_reference
    from oracle.ref_forward import ref_forward
    from oracle.synth import seeded_state_dict, synth_image
    from yolo_series_b200 import Model, _lib, plan
    path = f'/root/reference/cfg/{rel}'
    ours = Model(path)
    assert type(ours.model[-1]).__name__ == head
    RefModel, _ = import_reference()
    with torch.no_grad():
        ref = RefModel(path).eval()
        ref_sd = ref.state_dict()
        sd = seeded_state_dict(ours, seed=0)
        assert {k: tuple(v.shape) for k, v in ref_sd.items()} == {k: tuple(v.shape) for k, v in ours.state_dict().items()}
        assert ours.stride.tolist() == ref.stride.tolist()
        # lowering: concat / upsample planned away, aux branches of the P6 training cfgs dropped
        L = plan.lower(ours.model, in_ch=3)
        kinds = [op.kind for op in L.ops]
        assert _lib.OP_COPY not in kinds and _lib.OP_UPSAMPLE2 not in kinds
        n_conv_nodes = sum(1 for m in ref.modules() if type(m).__name__ in ('Conv', 'RepConv')) + ref.model[-1].nl
        assert 0 < kinds.count(_lib.OP_CONV) <= n_conv_nodes
        assert len(L.head_vals) == ref.model[-1].nl
        # oracle forward == real reference forward on this cfg (fused)
        ref.load_state_dict(sd)
        ref.fuse()
        s = int(ref.stride.max())
        x = synth_image(1, 2 * s, 2 * s, seed=2)
        want = ref(x)[0]
        with open(path) as f:
            cfg = yaml.safe_load(f)
        got, _ = ref_forward(cfg, sd, x, fused=True)
    assert got.shape == want.shape
    assert torch.allclose(got, want, rtol=2e-5, atol=2e-4), float((got - want).abs().max())
