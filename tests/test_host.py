"""CPU-side checks: cfg parser / module mirror / folding / lowering / C-ABI surface."""
import ctypes
import os
import re

import pytest
import torch
import torch.nn.functional as F

from oracle.ref_forward import ref_forward
from oracle.synth import seeded_state_dict, synth_image
from yolo_series_b200 import Model, _lib, fold, plan, zoo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, 'include', 'yolo_b200.h')).read()
    declared = set(re.findall(r'\b(yb_[a-z0-9_]+)\s*\(', hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert _lib.lib().yb_abi_version() == 4
    assert _lib.lib().yb_nms_workspace_bytes(2, 1000, 85, 0, 0) > 2 * 1000 * 16


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'yolo_series_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            src = open(os.path.join(pkg, fn)).read()
            assert 'import oracle' not in src and 'from oracle' not in src, fn


def test_model_surface_and_strides():
    m = Model('yolov7')
    assert m.stride.tolist() == [8., 16., 32.] and len(m.model) == 106
    h = m.model[-1]
    assert (h.nl, h.na, h.nc, h.no) == (3, 3, 80, 85)
    assert torch.allclose(h.anchors[0] * 8, h.anchor_grid[0].view(3, 2))
    assert next(m.parameters()).dtype == torch.float32
    w6 = Model('yolov7-w6-custom')
    assert w6.stride.tolist() == [8., 16., 32., 64.] and w6.model[-1].no == 88
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 3, 64, 64))       # no CPU fallback


def test_lowering_statistics():
    from collections import Counter
    # conv counts are after merging the sibling 1x1 convs that open each ELAN block (8 / 8 / 11 / 22 pairs)
    want = {'yolov7': (84, 5, 1, 0), 'yolov7-tiny': (50, 3, 1, 0), 'yolov7-w6-custom': (96, 0, 1, 0),
            'yolov7-e6e': (222, 8, 1, 11)}
    for name, (nconv, nmp, nspp, nadd) in want.items():
        L = plan.lower(Model(name).model, 3)
        c = Counter(o.kind for o in L.ops)
        assert (c[_lib.OP_CONV], c[_lib.OP_MAXPOOL2], c[_lib.OP_SPP3], c[_lib.OP_ADD]) == (nconv, nmp, nspp, nadd)
        assert c[_lib.OP_COPY] == 0 and c[_lib.OP_UPSAMPLE2] == 0     # concat + upsample fully planned away


def test_epilogue_fusion_pass_folds_pools_and_shortcuts(monkeypatch):
    """Engine._fuse_elementwise (host logic, no GPU): in the bf16 path every Shortcut add moves into the conv that is
    computed later (models/common.py:80-86) and every 2x2 max-pool behind a 1x1 conv into that conv (common.py:30-36);
    the fp32 parity mode and YB_NO_EPILOGUE_FUSION keep the lowered list; the shared lowered graph is not modified."""
    from collections import Counter

    import torch

    from yolo_series_b200.engine import Engine
    # (ops after fusion, pools fused, adds fused, pools left)
    want = {'yolov7': (85, 5, 0, 0), 'yolov7-tiny': (51, 3, 0, 0), 'yolov7-w6-custom': (97, 0, 0, 0),
            'yolov7-e6e': (224, 7, 11, 1)}
    for name, (nops, npool, nadd, pools_left) in want.items():
        L = plan.lower(Model(name).model, 3)
        before = list(L.ops)
        ops, pool_of = Engine._fuse_elementwise(L, torch.bfloat16)
        assert L.ops == before                                       # engines of one model share the lowered graph
        c = Counter(o.kind for o in ops)
        assert (len(ops), len(pool_of), c[_lib.OP_ADD], c[_lib.OP_MAXPOOL2]) == (nops, npool, 0, pools_left), name
        res = [o for o in ops if o.kind == _lib.OP_CONV and len(o.ins) == 2]
        assert len(res) == nadd
        for o in res:                                                # a residual conv is 1x1, bf16 out, no upsample
            assert o.conv.k == 1 and o.upsample == 1 and not L.vals[o.outs[0]].f32 and o.name.endswith('+add')
            assert L.vals[o.ins[1]].C == o.conv.c_out                # the other term has the conv's width
        made_at = {}                                                 # value id -> index of the op that writes it
        for i, o in enumerate(ops):
            for v in list(o.outs) + ([pool_of[id(o)]] if id(o) in pool_of else []):
                made_at[v] = i
        for i, o in enumerate(ops):                                  # the residual term is ready when its conv runs
            if o in res:
                assert made_at[o.ins[1]] < i, (name, o.name)
        for o in before:                                             # what a removed pool produced is still produced
            if o.kind == _lib.OP_MAXPOOL2 and o not in ops:
                assert o.outs[0] in pool_of.values()
        for cid, pv in pool_of.items():
            conv = next(o for o in ops if id(o) == cid)
            full = L.vals[conv.outs[0]]
            assert conv.conv.k == 1 and L.vals[pv].den == 2 * full.den and L.vals[pv].C == full.C
        ops32, pool32 = Engine._fuse_elementwise(L, torch.float32)
        assert ops32 == before and pool32 == {}
    L = plan.lower(Model('yolov7-e6e').model, 3)
    monkeypatch.setenv('YB_NO_EPILOGUE_FUSION', '1')
    ops, pool_of = Engine._fuse_elementwise(L, torch.bfloat16)
    assert len(ops) == len(L.ops) == 242 and pool_of == {}


def _run_lowered_cpu(L, x):
    """Interpret the lowered op list with torch CPU ops (test-only) to validate fold + placement logic."""
    vals = {}

    def root(v):
        off = 0
        while L.vals[v].parent is not None:
            off += L.vals[v].offset
            v = L.vals[v].parent
        return v, off
    bufs = {}

    def read(v):
        r, off = root(v)
        return bufs[r][:, off:off + L.vals[v].C]

    def write(v, t):
        r, off = root(v)
        if r not in bufs:
            bufs[r] = torch.zeros((t.shape[0], L.vals[r].C, t.shape[2], t.shape[3]))
        bufs[r][:, off:off + L.vals[v].C] = t
    if L.reorg:
        x = torch.cat([x[..., ::2, ::2], x[..., 1::2, ::2], x[..., ::2, 1::2], x[..., 1::2, 1::2]], 1)
    bufs[L.input_val] = x
    for op in L.ops:
        a = read(op.ins[0])
        if op.kind == _lib.OP_CONV:
            fc = op.conv
            y = F.conv2d(a, fc.weight, fc.bias, fc.s, fc.k // 2)
            y = F.silu(y) if fc.act == 1 else (F.leaky_relu(y, fc.slope) if fc.act == 2 else y)
            if op.upsample == 2:
                y = F.interpolate(y, scale_factor=2, mode='nearest')
            write(op.outs[0], y)
        elif op.kind == _lib.OP_MAXPOOL2:
            write(op.outs[0], F.max_pool2d(a, 2, 2))
        elif op.kind == _lib.OP_SPP3:
            for o, k in zip(op.outs, (5, 9, 13)):
                write(o, F.max_pool2d(a, k, 1, k // 2))
        elif op.kind == _lib.OP_POOL_S1:
            write(op.outs[0], F.max_pool2d(a, op.k, 1, op.k // 2))
        elif op.kind == _lib.OP_ADD:
            write(op.outs[0], a + read(op.ins[1]))
        elif op.kind == _lib.OP_COPY:
            write(op.outs[0], a)
        elif op.kind == _lib.OP_UPSAMPLE2:
            write(op.outs[0], F.interpolate(a, scale_factor=2, mode='nearest'))
    return [read(v) for v in L.head_vals]


@pytest.mark.parametrize('name,hw', [('yolov7-tiny', 64), ('yolov7', 64), ('yolov7-e6e', 128)])
def test_fold_and_placement_reproduce_oracle_logits(name, hw):
    """The lowered graph (folded weights, concat placement, fused upsample, dead aux removal) computes the
    same head logits as the oracle's node-by-node forward."""
    m = Model(name)
    sd = seeded_state_dict(m, seed=0)
    m.load_state_dict(sd)
    x = synth_image(1, hw, hw, seed=3)
    with torch.no_grad():
        L = plan.lower(m.model, 3)
        heads = _run_lowered_cpu(L, x)
        _, raws = ref_forward(zoo.get_cfg(name), sd, x, fused=True)
    for h, r in zip(heads, raws):
        b, na, ny, nx, no = r.shape
        got = h.view(b, na, no, ny, nx).permute(0, 1, 3, 4, 2)
        assert torch.allclose(got, r, rtol=1e-4, atol=1e-4), float((got - r).abs().max())


def test_repconv_identity_branch_fold():
    from yolo_series_b200.modules import RepConv
    torch.manual_seed(0)
    m = RepConv(16, 16, 3, 1)
    for bn in (m.rbr_identity, m.rbr_dense[1], m.rbr_1x1[1]):
        bn.running_mean.normal_(0, 0.2); bn.running_var.uniform_(0.5, 1.5); bn.weight.data.uniform_(0.5, 1.5); bn.bias.data.normal_()
        bn.eps = 1e-3
    x = torch.randn(1, 16, 8, 8)
    with torch.no_grad():
        def bnf(bn, t): return F.batch_norm(t, bn.running_mean, bn.running_var, bn.weight, bn.bias, False, 0., bn.eps)
        ref = F.silu(bnf(m.rbr_dense[1], F.conv2d(x, m.rbr_dense[0].weight, None, 1, 1)) +
                     bnf(m.rbr_1x1[1], F.conv2d(x, m.rbr_1x1[0].weight, None, 1, 0)) + bnf(m.rbr_identity, x))
        fc = fold.fold_repconv(m)
        got = F.silu(F.conv2d(x, fc.weight, fc.bias, 1, 1))
    assert torch.allclose(got, ref, atol=1e-5)


def test_tta_pass_geometry_matches_reference_shapes(gold):
    """Model._passes / rows_per_image reproduce scale_img's sizes (torch_utils.py:247-257) and the row count of the
    reference's augmented output (tests/golden/tta_tiny.npz, tta_scale_img.npz)."""
    m = Model('yolov7-tiny')
    assert m.rows_per_image(128, 160) == 1260
    assert m.rows_per_image(128, 160, augment=True) == gold('tta_tiny.npz')['pred'].shape[1] == 3276
    p = m._passes(96, 136, True)
    g = gold('tta_scale_img.npz')
    assert [q[0] for q in p] == [1, 0.83, 0.67] and [q[1] for q in p] == [None, 3, None]
    assert p[1][3] == g['r0.83_fNone'].shape[2:] and p[2][3] == g['r0.67_fNone'].shape[2:]
    assert p[1][2] == (int(96 * 0.83), int(136 * 0.83))
    w6 = Model('yolov7-w6-custom')                       # gs = 64 for the four-level heads
    assert [q[3] for q in w6._passes(1280, 1280, True)] == [(1280, 1280), (1088, 1088), (896, 896)]
    assert w6.rows_per_image(1280, 1280) == 102000


def test_ensemble_and_eval_surface_without_gpu():
    from yolo_series_b200 import Ensemble
    from yolo_series_b200 import evaluate
    ens = Ensemble([Model('yolov7-tiny'), Model('yolov7-tiny')])
    assert len(ens) == 2 and ens.stride.tolist() == [8., 16., 32.] and len(ens.names) == 80
    with pytest.raises(RuntimeError):
        ens(torch.zeros(1, 3, 64, 64))                   # no CPU fallback
    with pytest.raises(RuntimeError):
        evaluate.box_iou(torch.zeros(2, 4), torch.zeros(3, 4))
    with pytest.raises(RuntimeError):
        evaluate.match_batch(torch.zeros(1, 4, 6), torch.zeros(1, dtype=torch.int32), torch.zeros(0, 6), (64, 64),
                             [dict(shape0=(64, 64), ratio=(1, 1), pad=(0, 0))], torch.linspace(0.5, 0.95, 10))
    assert evaluate.NMS.conf == 0.25 and evaluate.NMS.iou == 0.45 and evaluate.autoShape.classes is None


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the reference algorithm on the host cores, the driver's denominator): ONE JSON line with
    the same metric string the b200 arm prints for that workload, `impl`, a `cpu_baseline` describing the run and an
    `e2e` object without copies.  Runs on the CPU (small cfg / size so that it takes seconds)."""
    import json
    import os
    import subprocess
    import sys

    import bench
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, 'bench.py'), '--impl', 'reference', '--model', 'yolov7-tiny',
                        '--img', '320', '--steps', '1', '--warmup', '0'], capture_output=True, text=True, timeout=600, cwd=root)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.strip().splitlines() if ln.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['metric'] == bench.metric_name('yolov7-tiny', 320)
    for k in ('value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling', 'vs_baseline',
              'dtype', 'data', 'config', 'cpu_baseline', 'e2e', 'gpu_launches'):
        assert k in d, k
    assert d['unit'] == 'images/s' and d['higher_is_better'] is True and d['vs_baseline'] is None and d['n_gpus'] == 1
    assert d['value'] > 0 and d['gpu_launches'] == 0 and 'workload' in d['config']
    cb, e = d['cpu_baseline'], d['e2e']
    assert cb['kind'] == 'port' and cb['cores'] >= 1 and cb['value'] == d['value'] and 'sample' in cb
    assert e['value'] == d['value'] and e['unit'] == d['unit'] and e['h2d_bytes_per_step'] == 0 and e['d2h_bytes_per_step'] == 0
