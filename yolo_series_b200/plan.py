"""Graph lowering: cfg node list (reference models/yolo.py:601-631 forward_once semantics) ->
flat list of primitive ops over NHWC tensor views, with the data-movement nodes planned away:

* ``Concat`` (common.py:61-62) never copies: each producer is *placed* at its channel offset inside
  the consumer's buffer and writes there directly (nested concats, e.g. DownC -> Concat, collapse).
  A producer feeding a second Concat is the only case that needs a copy op.
* ``nn.Upsample`` directly after a conv is folded into that conv's store (upsample=2).
* ``SPPCSPC`` / ``DownC`` / ``RepConv`` / head implicit layers are expanded / folded (fold.py).
* three ``SP`` nodes (5, 9, 13) reading one tensor (yolov7-tiny) become one fused ``spp3`` pass.
* the aux convs of the P6 training cfgs that ``IAuxDetect.fuseforward`` never reads
  are dead code and are dropped unless ``keep_aux=True``.

The result is independent of batch size / resolution except for shapes, which are resolved by
``instantiate``.
"""
from dataclasses import dataclass, field
from typing import List, Optional

from . import fold
from ._lib import (ACT_NONE, OP_ADD, OP_CONV, OP_COPY, OP_MAXPOOL2, OP_POOL_S1, OP_SPP3, OP_UPSAMPLE2)


@dataclass
class Val:
    """A logical NHWC tensor.  ``down`` = spatial reduction vs the network input (may be < 1)."""
    vid: int
    C: int
    num: int = 1          # H = Hin * num // den
    den: int = 1
    parent: Optional[int] = None   # concat value this one lives in
    offset: int = 0
    is_concat: bool = False
    f32: bool = False     # head logits are always fp32
    users: int = 0


@dataclass
class POp:
    kind: int
    ins: List[int]
    outs: List[int]
    conv: Optional[fold.FoldedConv] = None
    k: int = 0
    upsample: int = 1
    name: str = ''


@dataclass
class Lowered:
    vals: List[Val] = field(default_factory=list)
    ops: List[POp] = field(default_factory=list)
    input_val: int = 0
    reorg: bool = False
    head_vals: List[int] = field(default_factory=list)
    aux_vals: List[int] = field(default_factory=list)
    in_ch: int = 3


def _kind(m):
    return type(m).__name__


def lower(layers, in_ch=3, fused=False, keep_aux=False):
    """``layers``: the Sequential of nodes with ``.f`` / ``.i`` attributes (yolo.py:806)."""
    L = Lowered(in_ch=in_ch)

    def new(C, num, den, **kw):
        v = Val(len(L.vals), C, num, den, **kw)
        L.vals.append(v)
        return v.vid

    def emit(kind, ins, outs, **kw):
        L.ops.append(POp(kind, list(ins), list(outs), **kw))
        for i in ins:
            L.vals[i].users += 1

    def conv(fc, vin, name, f32=False):
        vi = L.vals[vin]
        assert fc.c_in == vi.C or (vin == L.input_val), f'{name}: Cin {fc.c_in} != {vi.C}'
        den = vi.den * fc.s
        out = new(fc.c_out, vi.num, den, f32=f32)
        emit(OP_CONV, [vin], [out], conv=fc, k=fc.k, name=name)
        return out

    def pool2(vin, name):
        vi = L.vals[vin]
        out = new(vi.C, vi.num, vi.den * 2)
        emit(OP_MAXPOOL2, [vin], [out], name=name)
        return out

    def concat(vins, name):
        out = new(sum(L.vals[v].C for v in vins), L.vals[vins[0]].num, L.vals[vins[0]].den, is_concat=True)
        off = 0
        for v in vins:
            vv = L.vals[v]
            if vv.parent is None and v != L.input_val:
                vv.parent, vv.offset = out, off          # producer writes in place
            else:                                         # already lives elsewhere -> explicit copy
                dst = new(vv.C, vv.num, vv.den, parent=out, offset=off)
                emit(OP_COPY, [v], [dst], name=name + '.copy')
            off += vv.C
        return out

    L.input_val = new(in_ch, 1, 1)
    outs = []          # value id per node index
    head = None
    sp_pending = {}    # input value -> {k: (node index)}
    for idx, m in enumerate(layers):
        f = m.f
        srcs = [outs[idx + j if j < 0 else j] for j in ([f] if isinstance(f, int) else f)] if idx or f != -1 else None
        if idx == 0:
            srcs = [L.input_val]
        kind, name = _kind(m), f'model.{idx}'
        if kind == 'Conv':
            out = conv(fold.fold_conv_block(m), srcs[0], name)
        elif kind == 'RepConv':
            out = conv(fold.fold_repconv(m), srcs[0], name)
        elif kind == 'Concat':
            out = concat(srcs, name)
        elif kind == 'MP':
            out = pool2(srcs[0], name)
        elif kind == 'SP':
            k = m.m.kernel_size if hasattr(m, 'm') else m.k
            vi = L.vals[srcs[0]]
            out = new(vi.C, vi.num, vi.den)
            emit(OP_POOL_S1, [srcs[0]], [out], k=int(k), name=name)
        elif kind in ('Upsample',):
            vi = L.vals[srcs[0]]
            out = new(vi.C, vi.num * 2, vi.den)
            emit(OP_UPSAMPLE2, [srcs[0]], [out], name=name)
        elif kind == 'ReOrg':
            assert idx == 0, 'ReOrg is only supported as the first node (yolov7-w6/e6/d6/e6e)'
            L.reorg = True
            L.vals[L.input_val].C = in_ch * 4
            L.vals[L.input_val].den = 2
            out = L.input_val
        elif kind == 'Shortcut':
            va = L.vals[srcs[0]]
            out = new(va.C, va.num, va.den)
            emit(OP_ADD, srcs[:2], [out], name=name)
        elif kind == 'DownC':
            a = conv(fold.fold_conv_block(m.cv2), conv(fold.fold_conv_block(m.cv1), srcs[0], name + '.cv1'), name + '.cv2')
            b = conv(fold.fold_conv_block(m.cv3), pool2(srcs[0], name + '.mp'), name + '.cv3')
            out = concat([a, b], name)
        elif kind == 'SPPCSPC':
            x1 = conv(fold.fold_conv_block(m.cv1), srcs[0], name + '.cv1')
            x1 = conv(fold.fold_conv_block(m.cv3), x1, name + '.cv3')
            x1 = conv(fold.fold_conv_block(m.cv4), x1, name + '.cv4')
            ks = [p.kernel_size if isinstance(p.kernel_size, int) else p.kernel_size[0] for p in m.m]
            vi = L.vals[x1]
            pooled = [new(vi.C, vi.num, vi.den) for _ in ks]
            if list(ks) == [5, 9, 13]:
                emit(OP_SPP3, [x1], pooled, name=name + '.m')
            else:
                for kk, pv in zip(ks, pooled):
                    emit(OP_POOL_S1, [x1], [pv], k=int(kk), name=name + f'.m{kk}')
            cat1 = concat([x1] + pooled, name + '.cat1')
            y1 = conv(fold.fold_conv_block(m.cv5), cat1, name + '.cv5')
            y1 = conv(fold.fold_conv_block(m.cv6), y1, name + '.cv6')
            y2 = conv(fold.fold_conv_block(m.cv2), srcs[0], name + '.cv2')
            out = conv(fold.fold_conv_block(m.cv7), concat([y1, y2], name + '.cat2'), name + '.cv7')
        elif kind in ('Detect', 'IDetect', 'IAuxDetect'):
            head = m
            for i, fc in enumerate(fold.fold_head(m, fused)):
                L.head_vals.append(conv(fc, srcs[i], f'{name}.m.{i}', f32=True))
            L.aux_vals = list(srcs[m.nl:]) if keep_aux else []
            out = None
        else:
            raise NotImplementedError(f'node {idx}: module {kind} is outside the yolov7 hot path (SURVEY.md 8a)')
        outs.append(out)
    if head is None:
        raise ValueError('model has no Detect/IDetect/IAuxDetect head')
    _fuse_sp_triples(L)
    _fuse_upsample(L)
    _drop_dead(L)
    _merge_sibling_convs(L)
    return L


def _fuse_sp_triples(L):
    """SP5/SP9/SP13 on one input (yolov7-tiny.yaml nodes 31-33) -> one spp3 op."""
    by_in = {}
    for op in L.ops:
        if op.kind == OP_POOL_S1:
            by_in.setdefault(op.ins[0], {})[op.k] = op
    for vin, d in by_in.items():
        if all(k in d for k in (5, 9, 13)):
            first = min(L.ops.index(d[k]) for k in (5, 9, 13))
            fused = POp(OP_SPP3, [vin], [d[5].outs[0], d[9].outs[0], d[13].outs[0]], name=d[5].name + '+spp3')
            L.vals[vin].users -= 2
            for k in (5, 9, 13):
                L.ops.remove(d[k])
            L.ops.insert(first, fused)


def _fuse_upsample(L):
    """conv -> Upsample where the conv has no other reader: the conv stores the 2x2 replicas itself."""
    for up in [o for o in L.ops if o.kind == OP_UPSAMPLE2]:
        src = up.ins[0]
        prod = next((o for o in L.ops if src in o.outs), None)
        if prod is None or prod.kind != OP_CONV or L.vals[src].users != 1 or prod.upsample != 1:
            continue
        if L.vals[src].parent is not None:
            continue
        prod.outs = [up.outs[0]]
        prod.upsample = 2
        prod.name += '+up2'
        L.ops.remove(up)


def _merge_sibling_convs(L):
    """Two 1x1/s1 convs that read the same tensor and write neighbouring channel slices of one concat buffer
    (the cv1 / cv2 pair that opens every ELAN block, cfg/deploy/yolov7.yaml nodes 4-5, 17-18, ...) become one
    conv with the filters stacked: the input is read once and one kernel is launched.  Per output channel the
    arithmetic is unchanged, so results are bit-identical."""
    import torch
    i = 0
    while i < len(L.ops):
        a = L.ops[i]
        i += 1
        if a.kind != OP_CONV or a.conv.k != 1 or a.conv.s != 1 or a.upsample != 1:
            continue
        va = L.vals[a.outs[0]]
        if va.parent is None or va.f32:
            continue
        for b in L.ops[i:]:
            if b.kind != OP_CONV or b.ins != a.ins or b.conv.k != 1 or b.conv.s != 1 or b.upsample != 1:
                continue
            vb = L.vals[b.outs[0]]
            if vb.parent != va.parent or vb.f32 or (b.conv.act, b.conv.slope) != (a.conv.act, a.conv.slope):
                continue
            if va.offset + va.C == vb.offset:
                lo, hi, vlo = a, b, va
            elif vb.offset + vb.C == va.offset:
                lo, hi, vlo = b, a, vb
            else:
                continue
            fc = fold.FoldedConv(torch.cat([lo.conv.weight, hi.conv.weight], 0), torch.cat([lo.conv.bias, hi.conv.bias], 0),
                                 1, 1, a.conv.act, a.conv.slope)
            merged = Val(len(L.vals), va.C + vb.C, va.num, va.den, parent=va.parent, offset=vlo.offset)
            L.vals.append(merged)
            a.conv, a.outs, a.name = fc, [merged.vid], lo.name + '+' + hi.name
            L.vals[a.ins[0]].users -= 1
            L.ops.remove(b)
            break


def _drop_dead(L):
    """Remove ops whose outputs nobody reads (aux branches of the P6 cfgs)."""
    live = set(L.head_vals) | set(L.aux_vals)

    def root(v):
        while L.vals[v].parent is not None:
            v = L.vals[v].parent
        return v

    changed = True
    while changed:
        changed = False
        needed = set(live)
        for op in L.ops:
            for i in op.ins:
                needed.add(i)
        # a value placed inside a concat is needed iff the concat (root) is needed
        roots_needed = {root(v) for v in needed} | needed
        keep = []
        for op in L.ops:
            if any((o in needed) or (L.vals[o].parent is not None and root(o) in roots_needed) for o in op.outs):
                keep.append(op)
            else:
                changed = True
        L.ops = keep
