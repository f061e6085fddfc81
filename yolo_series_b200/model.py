"""Drop-in ``Model`` for the reference's ``models.yolo.Model`` (models/yolo.py:508-733), running the
forward pass on the B200 engine.

Same constructor ``Model(cfg, ch=3, nc=None, anchors=None)`` (yolo.py:509), same call
``model(x, augment=False, profile=False)`` (yolo.py:581) returning the eval tuple
``(pred (B, A, no), [raw_i (B, na, ny, nx, no)])`` (yolo.py:174), same attributes callers read
(``stride``, ``names``, ``yaml``, ``model``, ``save``, ``model[-1].nl/na/nc/no/anchors/anchor_grid``,
``fuse()``, ``half()/float()/to()/eval()``, ``parameters()``) -- see SURVEY.md 8(b).

``accelerate(ref_model)`` wraps an already-built reference ``Model`` (e.g. unpickled by
``attempt_load``, models/experimental.py:247-270) without re-creating it.
"""
import ctypes as C
import math
import os
from collections import OrderedDict
from copy import deepcopy

import torch
import torch.nn as nn

from . import _lib, modules, plan, zoo
from .engine import Engine


def make_divisible(x, divisor):
    return math.ceil(x / divisor) * divisor


def _resolve_arg(a, scope):
    """cfg strings are python expressions in the reference (yolo.py:744-749): support the ones the
    yolov7 cfgs use without eval()."""
    if not isinstance(a, str):
        return a
    if a in scope:
        return scope[a]
    if a == 'None':
        return None
    if a.startswith('nn.LeakyReLU'):
        inner = a[a.index('(') + 1:a.rindex(')')].strip()
        return nn.LeakyReLU(float(inner) if inner else 0.01)
    if a in ('nn.SiLU()', 'nn.SiLU'):
        return nn.SiLU()
    if a in ('nn.Identity()',):
        return nn.Identity()
    return a  # e.g. 'nearest'


def _sources(spec, index):
    """Absolute producer indices of a cfg row's ``from`` field (negative = relative to ``index``)."""
    refs = [spec] if isinstance(spec, int) else list(spec)
    return [index + r if r < 0 else r for r in refs]


# Channel rule per module family: (input widths, resolved args, head width, width multiple) -> (ctor args, Cout).
# ``None`` as Cout means "no feature map" (the detection head ends the graph).
def _rule_conv(cin, args, head_no, gw, repeats=None):
    cout = args[0] if args[0] == head_no else make_divisible(args[0] * gw, 8)
    ctor = [cin[0], cout] + ([] if repeats is None else [repeats]) + list(args[1:])
    return ctor, cout


def _rule_head(cin, args, head_no, gw):
    ctor = list(args)
    if isinstance(ctor[1], int):                       # anchors given as a count: placeholder values per level
        ctor[1] = [list(range(ctor[1] * 2))] * len(cin)
    return ctor + [list(cin)], None


_CHANNEL_RULES = {
    'Conv': _rule_conv, 'RepConv': _rule_conv,
    'DownC': _rule_conv, 'SPPCSPC': _rule_conv,          # these two take the repeat count as their third argument
    'Concat': lambda c, a, h, g: (list(a), sum(c)),
    'Shortcut': lambda c, a, h, g: (list(a), c[0]),
    'ReOrg': lambda c, a, h, g: (list(a), 4 * c[0]),
    'Detect': _rule_head, 'IDetect': _rule_head, 'IAuxDetect': _rule_head,
}
_PASS_THROUGH = lambda c, a, h, g: (list(a), c[0])     # MP, SP, nn.Upsample: width unchanged


def parse_model(d, ch):
    """cfg dict -> (nn.Sequential of parameter containers with .i/.f/.type/.np, save list).
    Same grammar as the reference's builder (yolo.py:736-813), driven by ``_CHANNEL_RULES``
    instead of an if-chain; ``eval()`` of cfg strings is replaced by ``_resolve_arg`` + ``modules.REGISTRY``."""
    anchors = d['anchors']
    per_level = (len(anchors[0]) // 2) if isinstance(anchors, list) else anchors
    head_no = per_level * (d['nc'] + 5)
    gd, gw = d['depth_multiple'], d['width_multiple']
    names = {'nc': d['nc'], 'anchors': anchors}
    rows = list(d['backbone']) + list(d['head'])
    widths, nodes, kept = [], [], set()
    for index, (spec, repeats, mname, raw_args) in enumerate(rows):
        cls = modules.REGISTRY.get(mname)
        if cls is None:
            raise NotImplementedError(f'cfg node {index}: module {mname!r} is outside the yolov7 hot path')
        depth = max(round(repeats * gd), 1) if repeats > 1 else repeats
        takes_depth = mname in ('DownC', 'SPPCSPC')
        if depth != 1 and not takes_depth:
            raise NotImplementedError('depth repeats > 1 do not occur in the yolov7 cfgs')
        srcs = _sources(spec, index)
        cin = [ch[-1]] if index == 0 else [widths[j] for j in srcs]
        rule = _CHANNEL_RULES.get(mname, _PASS_THROUGH)
        resolved = [_resolve_arg(a, names) for a in raw_args]
        ctor, cout = rule(cin, resolved, head_no, gw, depth) if takes_depth else rule(cin, resolved, head_no, gw)
        node = cls(*ctor)
        node.i, node.f, node.type = index, spec, mname
        node.np = sum(p.numel() for p in node.parameters())
        kept.update(j for j, r in zip(srcs, [spec] if isinstance(spec, int) else spec) if r != -1)
        nodes.append(node)
        widths.append(cout)
    return nn.Sequential(*nodes), sorted(kept)


def _node_strides(layers):
    """Spatial reduction of every node output vs the input (replaces the 256x256 dry run of
    yolo.py:535-536)."""
    red = []
    for idx, m in enumerate(layers):
        f = m.f
        srcs = [idx + j if j < 0 else j for j in ([f] if isinstance(f, int) else f)]
        r_in = 1.0 if idx == 0 else red[srcs[0]]
        kind = type(m).__name__
        if kind in ('Conv', 'RepConv'):
            conv = m.conv if hasattr(m, 'conv') else (m.rbr_reparam if hasattr(m, 'rbr_reparam') else m.rbr_dense[0])
            r = r_in * conv.stride[0]
        elif kind in ('MP', 'DownC', 'ReOrg'):
            r = r_in * 2
        elif kind == 'Upsample':
            r = r_in / 2
        else:
            r = r_in
        red.append(r)
    return red


class Model(nn.Module):
    def __init__(self, cfg='yolov7', ch=3, nc=None, anchors=None):
        super().__init__()
        self.traced = False
        if isinstance(cfg, dict):
            self.yaml = deepcopy(cfg)
        elif isinstance(cfg, str) and os.path.exists(cfg):
            import yaml
            self.yaml_file = os.path.basename(cfg)
            with open(cfg) as f:
                self.yaml = yaml.safe_load(f)
        else:
            self.yaml = zoo.get_cfg(cfg)
            self.yaml_file = str(cfg)
        ch = self.yaml['ch'] = self.yaml.get('ch', ch)
        if nc and nc != self.yaml['nc']:
            self.yaml['nc'] = nc
        if anchors:
            self.yaml['anchors'] = round(anchors)
        self.model, self.save = parse_model(deepcopy(self.yaml), ch=[ch])
        self.names = [str(i) for i in range(self.yaml['nc'])]
        head = self.model[-1]
        if not isinstance(head, modules.HEADS):
            raise NotImplementedError('last node must be Detect / IDetect / IAuxDetect')
        red = _node_strides(self.model)
        srcs = [len(self.model) - 1 + j if j < 0 else j for j in head.f]
        head.stride = torch.tensor([float(red[s]) for s in srcs[:head.nl]])
        self._check_anchor_order(head)
        head.anchors /= head.stride.view(-1, 1, 1)
        self.stride = head.stride
        self._init_head_biases(head)
        for m in self.modules():
            if isinstance(m, nn.BatchNorm2d):     # utils/torch_utils.py:150-151
                m.eps, m.momentum = 1e-3, 0.03
        self._fused = False
        self._engines, self._packed = OrderedDict(), {}
        self._compute_dtype = torch.bfloat16
        self.use_graph = True

    # ---- construction helpers (yolo.py:633-655, utils/autoanchor.py:12-20)
    @staticmethod
    def _check_anchor_order(m):
        a = m.anchor_grid.prod(-1).view(-1)
        if (a[-1] - a[0]).sign() != (m.stride[-1] - m.stride[0]).sign():
            m.anchors[:] = m.anchors.flip(0)
            m.anchor_grid[:] = m.anchor_grid.flip(0)

    @staticmethod
    def _init_head_biases(m):
        convs = list(zip(m.m, m.stride)) + (list(zip(m.m2, m.stride)) if hasattr(m, 'm2') else [])
        with torch.no_grad():
            for mi, s in convs:
                b = mi.bias.view(m.na, -1)
                b[:, 4] += math.log(8 / (640 / float(s)) ** 2)
                b[:, 5:] += math.log(0.6 / (m.nc - 0.99))

    # ---- reference-compatible surface
    def fuse(self):
        """Folding happens when the engine is built; this only records the request (yolo.py:693-710)."""
        self._invalidate()
        return self

    def half(self):
        self._compute_dtype = torch.bfloat16   # fp16 on the reference (detect.py:49); bf16 here
        self._invalidate()
        return self

    def float(self):
        self._compute_dtype = torch.float32
        self._invalidate()
        return self

    def set_precision(self, mode):
        """'bf16' (tcgen05 tensor cores, default) or 'fp32' (CUDA-core reference mode)."""
        self._compute_dtype = {'bf16': torch.bfloat16, 'fp32': torch.float32}[mode]
        self._invalidate()
        return self

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        self._invalidate()
        return out

    # Engines are shape-specialised (buffers + TMA plans + CUDA graph: 5.5 GB at yolov7 / 640 / B=32) and the reference's
    # callers produce many shapes (test.py rect batches, detect.py streams, the short last batch, 3 TTA scales), so
    # the cache is a small LRU; evicted engines are destroyed at once.  Packed weights are shape-independent and are
    # shared by all engines of the model (`_packed`).
    engine_cache_size = 6

    def _invalidate(self):
        for e in getattr(self, '_engines', {}).values():
            e.destroy()
        self._engines = OrderedDict()
        self._packed = {}
        self._lowered = None

    def info(self, verbose=False, img_size=640):
        n = sum(p.numel() for p in self.parameters())
        print(f'Model Summary: {len(list(self.modules()))} layers, {n} parameters')

    # ---- engine plumbing
    def _layers(self):
        return self.model

    def _is_ref_fused(self):
        return False

    def _get_engine(self, B, H, W, device, img_dtype=torch.uint8):
        if device.index is None:
            device = torch.device('cuda', torch.cuda.current_device())
        # an engine owns its activation buffers, so calls issued on different streams must not share one: the key
        # carries the current stream (two batches in flight = two engines; the packed weights are shared)
        key = (B, H, W, str(device), self._compute_dtype, img_dtype, torch.cuda.current_stream(device).cuda_stream)
        eng = self._engines.get(key)
        if eng is None:
            if getattr(self, '_lowered', None) is None:
                with torch.no_grad():
                    self._lowered = plan.lower(self._layers(), in_ch=self.yaml.get('ch', 3),
                                               fused=self._is_ref_fused())
            while len(self._engines) >= max(1, int(self.engine_cache_size)):
                _, old = self._engines.popitem(last=False)       # least recently used
                torch.cuda.synchronize(device)                   # it may still be running on another stream
                old.destroy()
            if not hasattr(self, '_packed'):
                self._packed = {}
            eng = Engine(self._lowered, B, H, W, device, self._compute_dtype, self.use_graph,
                         getattr(self, 'tile_overrides', None), img_dtype=img_dtype,
                         fuse_stem=getattr(self, 'fuse_stem', True), packed=self._packed)
            self._engines[key] = eng
        else:
            self._engines.move_to_end(key)
        return eng

    def _device(self):
        p = next(self.parameters())
        return p.device

    def compute_streams(self, n, device=None):
        """``n`` side streams of ``device``, created once per model (engines are cached per stream, so the streams have
        to be the same objects from call to call)."""
        dev = device if device is not None else self._device()
        if dev.index is None:
            dev = torch.device('cuda', torch.cuda.current_device())
        pool = self.__dict__.setdefault('_side_streams', {}).setdefault(str(dev), [])
        while len(pool) < n:
            pool.append(torch.cuda.Stream(dev))
        return pool[:n]

    def head_levels(self, eng):
        head = self._layers()[-1]
        cached = getattr(eng, '_head_lv', None)    # anchors are read back once per engine: .cpu() would drain the stream
        if cached is not None:
            return cached, head
        lv = (_lib.HeadLevel * head.nl)()
        ag = head.anchor_grid.detach().float().cpu().view(head.nl, -1, 2)
        for i in range(head.nl):
            lv[i].logits = eng.head_views[i]
            lv[i].stride = float(head.stride[i])
            for a in range(head.na):
                lv[i].anchor_wh[2 * a] = float(ag[i, a, 0])
                lv[i].anchor_wh[2 * a + 1] = float(ag[i, a, 1])
        eng._head_lv = lv
        return lv, head

    def run_trunk(self, x):
        """Stage + trunk + head convs.  Returns the engine (its ``head`` buffers hold the logits)."""
        dev = x.device
        if dev.type != 'cuda':
            dev = self._device()
            if dev.type != 'cuda':
                raise RuntimeError('yolo_series_b200.Model needs a CUDA device: move the model or the input '
                                   'to cuda (no CPU fallback)')
        if x.dtype not in (torch.uint8, torch.float32):
            x = x.float()
        B, _, H, W = x.shape
        gs = int(self.stride.max())
        if H % gs or W % gs:
            raise ValueError(f'input {H}x{W} must be a multiple of the max stride {gs} (general.py:124-129)')
        eng = self._get_engine(B, H, W, dev, x.dtype)
        eng.stage(x)
        eng.run()
        return eng

    # ---- TTA / ensemble plumbing (yolo.py:582-597, experimental.py:74-81)
    TTA_SCALES = (1, 0.83, 0.67)
    TTA_FLIPS = (None, 3, None)      # 2 = up-down, 3 = left-right

    def _passes(self, H, W, augment):
        """[(scale, flip, (h, w) resized, (Hp, Wp) padded)] of one call (scale_img, torch_utils.py:247-257)."""
        if not augment:
            return [(1, None, (H, W), (H, W))]
        gs = int(self.stride.max())
        out = []
        for si, fi in zip(self.TTA_SCALES, self.TTA_FLIPS):
            if si == 1.0:
                out.append((si, fi, (H, W), (H, W)))
            else:
                out.append((si, fi, (int(H * si), int(W * si)),
                            (math.ceil(H * si / gs) * gs, math.ceil(W * si / gs) * gs)))
        return out

    def rows_per_image(self, H, W, augment=False):
        """A of ``pred`` for an (H, W) input: na * sum of the level grids, summed over the TTA passes."""
        head = self._layers()[-1]
        strides = [int(s) for s in head.stride.tolist()]
        return sum(head.na * sum((hp // s) * (wp // s) for s in strides) for _, _, _, (hp, wp) in self._passes(H, W, augment))

    def _scaled_input(self, x, si, fi, hw, HW):
        """scale_img(x.flip(fi), si) as one kernel (yb_scale_img): fp32 (B, C, Hp, Wp) on the device."""
        dev = x.device if x.device.type == 'cuda' else self._device()
        if x.device != dev:
            x = x.to(dev, non_blocking=True)
        x = x.contiguous()
        B, Cc, H, W = x.shape
        out = torch.empty((B, Cc, HW[0], HW[1]), dtype=torch.float32, device=dev)
        code = _lib.YB_U8 if x.dtype == torch.uint8 else _lib.YB_F32
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().yb_scale_img(C.c_void_p(x.data_ptr()), code, B, Cc, H, W, int(fi or 0), hw[0], hw[1],
                                               C.c_void_p(out.data_ptr()), HW[0], HW[1], 0.447, _lib.stream_ptr()),
                       'yb_scale_img')
        return out

    def forward_into(self, x, pred, row_off, augment=False):
        """Decode this model's prediction rows for ``x`` into ``pred[:, row_off:row_off + rows]`` (B, rows_total, no)
        and return the number of rows written.  Used by ``forward(augment=True)`` and ``Ensemble`` so the
        ``torch.cat(y, 1)`` of yolo.py:597 / experimental.py:80 never happens as a copy."""
        if x.dtype not in (torch.uint8, torch.float32):
            x = x.float()
        B, _, H, W = x.shape
        written = 0
        for si, fi, hw, HW in self._passes(H, W, augment):
            xi = x if (si == 1.0 and not fi) else self._scaled_input(x, si, fi, hw, HW)
            eng = self.run_trunk(xi)
            lv, head = self.head_levels(eng)
            A = sum(head.na * v.H * v.W for v in eng.head_views)
            tp = _lib.TtaPass(pred.shape[1], row_off + written, float(si), int(fi or 0), float(H), float(W))
            with torch.cuda.device(eng.device):
                _lib.check(eng.lib.yb_decode_tta(lv, head.nl, head.na, head.no, C.c_void_p(pred.data_ptr()), C.byref(tp),
                                                 _lib.stream_ptr()), 'yb_decode_tta')
            written += A
        return written

    def _print_profile(self, eng):
        """``model(x, profile=True)`` (yolo.py:614-623, 629-630): the reference prints GFLOPS / params / ms per node and
        the total; here the unit of execution is the lowered op (one kernel launch), timed with CUDA events."""
        rows = eng.profile_table(3)
        print(f"{'ms':>9s} {'GFLOP':>9s} {'TFLOP/s':>8s} {'GB/s':>7s}  {'kernel':<16s} op")
        for r in rows:
            print(f"{r['ms']:9.4f} {r['flops'] / 1e9:9.2f} {r['flops'] / r['ms'] / 1e9 if r['flops'] else 0:8.1f} "
                  f"{r['bytes'] / r['ms'] / 1e6:7.0f}  {r['kernel']:<16s} {r['op']}")
        print(f"{sum(r['ms'] for r in rows):.3f} ms total (eager, one event per op), "
              f"{sum(r['flops'] for r in rows) / 1e9:.1f} GFLOP, batch {eng.B}")

    @torch.no_grad()
    def forward(self, x, augment=False, profile=False):
        if augment:
            # augmented inference (yolo.py:582-597): three passes decoded straight into one pred, de-scaled / de-flipped
            B, _, H, W = x.shape
            head = self._layers()[-1]
            dev = x.device if x.device.type == 'cuda' else self._device()
            pred = torch.empty((B, self.rows_per_image(H, W, True), head.no), dtype=torch.float32, device=dev)
            self.forward_into(x, pred, 0, augment=True)
            return pred, None
        eng = self.run_trunk(x)
        lv, head = self.head_levels(eng)
        A = sum(head.na * v.H * v.W for v in eng.head_views)
        pred = torch.empty((eng.B, A, head.no), dtype=torch.float32, device=eng.device)
        raws = [torch.empty((eng.B, head.na, v.H, v.W, head.no), dtype=torch.float32, device=eng.device)
                for v in eng.head_views]
        raw_ptrs = (C.c_void_p * head.nl)(*[r.data_ptr() for r in raws])
        with torch.cuda.device(eng.device):
            _lib.check(eng.lib.yb_decode(lv, head.nl, head.na, head.no, C.c_void_p(pred.data_ptr()), raw_ptrs,
                                         _lib.stream_ptr()), 'yb_decode')
        if profile:
            self._print_profile(eng)
        return pred, raws

    @torch.no_grad()
    def detect(self, x, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False, multi_label=False,
               max_det=300, max_nms=30000, packed=False, augment=False, gather=None):
        """Fused fast path: forward + decode + NMS without materialising ``pred`` (yb_decode_nms).
        Returns the same list as ``non_max_suppression(model(x)[0], ...)``; ``packed=True`` returns the device
        buffers ``(out (B,max_det,6), meta int32 [counts..., overflow])`` with no host synchronisation.
        ``gather``: a ``dist.PeerGather`` -- image-sharded multi-GPU inference: the NMS kernel stores every kept row into
        all GPUs' receive buffers over NVLink and the call returns ``(out_all (world*B, max_det, 6), meta_all)`` in
        global image order (``meta_all`` = the world*B counts followed by this rank's overflow flag)."""
        from .nms import _run_nms
        if augment and gather is not None:
            raise NotImplementedError('detect(augment=True, gather=...): the peer gather is fused into the single-pass '
                                      'decode+NMS kernel; gather TTA results with dist.gather_detections')
        if augment:      # TTA: the passes share one pred, then the stand-alone NMS (3x the candidates)
            pred = self.forward(x, augment=True)[0]
            return _run_nms(pred, pred.shape[0], pred.shape[1], pred.shape[2], conf_thres, iou_thres, classes, agnostic,
                            multi_label, max_nms, max_det, pred.device, return_packed=packed)
        eng = self.run_trunk(x)
        lv, head = self.head_levels(eng)
        A = sum(head.na * v.H * v.W for v in eng.head_views)
        if gather is not None:
            if getattr(eng, '_gather_warmed', None) is not gather:
                # first use of this engine: lowering, weight packing, TMA plans and graph capture just happened inside
                # run_trunk and take seconds -- line the ranks up before the first exchange so that nobody's bounded
                # wait starts while a peer is still building
                import torch.distributed as tdist
                torch.cuda.synchronize(eng.device)
                tdist.barrier(group=gather.group)
                eng._gather_warmed = gather
            table, slot, step = gather.begin(eng.B, max_det)
            out, meta = _run_nms(None, eng.B, A, head.no, conf_thres, iou_thres, classes, agnostic, multi_label, max_nms,
                                 max_det, eng.device, levels=(lv, head.nl, head.na), return_packed=True, peer=table)
            out_all, cnt_all = gather.finish(slot, step)
            meta_all = torch.cat((cnt_all, meta[-1:]))
            if packed:
                return out_all, meta_all
            host = meta_all.cpu()
            gather.raise_if_poisoned(host[:-1])
            if int(host[-1]) != 0:
                raise RuntimeError('yb_nms: candidate capacity exceeded (YB_ERR_OVERFLOW)')
            return [out_all[i, :int(host[i])] for i in range(out_all.shape[0])]
        return _run_nms(None, eng.B, A, head.no, conf_thres, iou_thres, classes, agnostic, multi_label, max_nms,
                        max_det, eng.device, levels=(lv, head.nl, head.na), return_packed=packed)


    @torch.no_grad()
    def detect_batches(self, batches, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False,
                       multi_label=False, max_det=300, max_nms=30000, post=None, gather=None, compute_streams=None):
        """Streaming form of ``detect`` for host batches (the detect.py:87-152 loop over a dataset): a generator that
        takes (B, 3, H, W) uint8/float32 HOST tensors (pinned for real overlap) and yields, in order, one
        ``(out (B', max_det, 6) fp32, counts (B',) int32)`` pair of pinned HOST tensors per batch.

        Copy streams and slots: batch i+1's H2D copy and batch i-1's D2H read run on their own copy streams
        while batch i computes, so PCIe time leaves the critical path; each step still moves its own input H2D and its
        own result D2H.  ``compute_streams`` = 2 (the default without ``post`` / with no or two ``gather`` objects)
        alternates the batches between two side streams with one engine each and six slots (3 or 4 streams: 9 / 12 slots;
        they only pay at very small batches -- batch 1: 0.49 ms per image with two, 0.39 with three): consecutive batches are
        independent, so the CTAs of batch i+1 fill the SMs that batch i's persistent kernels leave idle in their last
        round and between dependent launches -- +2.4 % images/s at batch 32 (`scripts/e2e_streams_probe.py`; two or
        three slots are too shallow: the host then enqueues a batch only when both streams are about to drain), ~+60 %
        at batch 1 (`scripts/two_stream_probe.py`).  1 computes on the caller's stream with two slots.  A
        ``dist.PeerGather``'s slot protocol is ordered by ONE stream: pass one gather per compute stream.
        ``post(out, counts) -> (out, counts)`` runs on the compute stream before the read-back (e.g.
        ``dist.gather_detections``); ``gather`` (a ``dist.PeerGather``, or one per compute stream) does the multi-GPU
        exchange inside the NMS kernel instead and every yielded pair then covers the global batch.  A yielded pair
        stays valid until the next batch is drawn (its pinned buffers are reused once every slot has been used again: 2 batches later with one compute stream, 3 per stream otherwise)."""
        from collections import deque
        dev = self._device()
        if dev.type != 'cuda':
            raise RuntimeError('yolo_series_b200.Model needs a CUDA device (no CPU fallback)')
        main = torch.cuda.current_stream(dev)
        gathers = list(gather) if isinstance(gather, (list, tuple)) else [gather]
        if compute_streams is None:      # a PeerGather's slot protocol is ordered by ONE stream: one gather per stream
            compute_streams = 2 if (post is None and (gather is None or len(gathers) == 2)) else 1
        if compute_streams not in (1, 2, 3, 4) or (gathers[0] is not None and len(gathers) < compute_streams):
            raise ValueError('compute_streams must be 1..4, with one PeerGather per compute stream')
        cstreams = [main] if compute_streams == 1 else self.compute_streams(compute_streams, dev)
        # copy streams, staging buffers and pinned result buffers live with the model: pinning host memory costs
        # milliseconds, which a fresh set per call would put into every short run (one generator per model at a time)
        cache = self.__dict__.setdefault('_stream_slots', {})
        key = (str(dev), compute_streams)
        if key not in cache:
            cache[key] = dict(h2d=torch.cuda.Stream(dev), d2h=torch.cuda.Stream(dev),
                              slots=[dict(stage=None, host=None, consumed=torch.cuda.Event(), loaded=torch.cuda.Event(),
                                          done=torch.cuda.Event(), read=torch.cuda.Event(), used=False, gather=None)
                                     for _ in range(2 if compute_streams == 1 else 3 * compute_streams)])
        h2d, d2h, slots = cache[key]['h2d'], cache[key]['d2h'], cache[key]['slots']
        for s in slots:                      # a previous generator that was abandoned half way may have left work behind
            if s['used']:
                s['read'].synchronize()
            s['keep'] = None
        pending = deque()

        def finish(s):
            s['read'].synchronize()
            meta = s['host'][1]
            B = meta.numel() - 1
            if int(meta[B]) != 0:
                raise RuntimeError('yb_nms: candidate capacity exceeded (YB_ERR_OVERFLOW)')
            if s['gather'] is not None:
                s['gather'].raise_if_poisoned(meta[:B])
            s['keep'] = None
            return s['host'][0], meta[:B]

        for i, xb in enumerate(batches):
            s = slots[i % len(slots)]
            if len(pending) == len(slots):
                yield finish(pending.popleft())
            if s['stage'] is None or s['stage'].shape != xb.shape or s['stage'].dtype != xb.dtype:
                s['stage'] = torch.empty(xb.shape, dtype=xb.dtype, device=dev)
            if s['used']:
                h2d.wait_event(s['consumed'])
            with torch.cuda.stream(h2d):
                s['stage'].copy_(xb, non_blocking=True)
                s['loaded'].record(h2d)
            cs = cstreams[i % len(cstreams)]
            g = gathers[i % len(cstreams)] if len(gathers) > 1 else gathers[0]
            cs.wait_event(s['loaded'])
            if s['used']:
                cs.wait_event(s['read'])         # the slot's previous result has left the device
            with torch.cuda.stream(cs):
                out, meta = self.detect(s['stage'], conf_thres, iou_thres, classes, agnostic, multi_label, max_det,
                                        max_nms, packed=True, gather=g)
                s['consumed'].record(cs)
                ovf = meta[-1:]
                if post is not None:
                    out, cnt = post(out, meta[:-1])
                    meta = torch.cat((cnt.to(torch.int32), ovf))
                s['done'].record(cs)
            if s['host'] is None or s['host'][0].shape != out.shape:
                s['host'] = (torch.empty(out.shape, dtype=out.dtype).pin_memory(),
                             torch.empty(meta.shape, dtype=meta.dtype).pin_memory())
            d2h.wait_event(s['done'])
            with torch.cuda.stream(d2h):
                s['host'][0].copy_(out, non_blocking=True)
                s['host'][1].copy_(meta, non_blocking=True)
                s['read'].record(d2h)
            s['keep'] = (out, meta)          # device results stay referenced until the read-back has completed
            s['used'], s['gather'] = True, g
            pending.append(s)
        while pending:
            yield finish(pending.popleft())


class AcceleratedModel(Model):
    """Wraps an existing reference ``Model`` instance (duck-typed) -- see ``accelerate``."""

    def __init__(self, ref_model):
        nn.Module.__init__(self)
        self.ref = ref_model
        self.traced = False
        self.yaml = getattr(ref_model, 'yaml', {'ch': 3})
        self.model = ref_model.model
        self.save = getattr(ref_model, 'save', [])
        self.names = getattr(ref_model, 'names', None)
        self.stride = ref_model.stride
        self._engines, self._packed, self._lowered = OrderedDict(), {}, None
        self._compute_dtype = torch.bfloat16
        self.use_graph = True

    def _is_ref_fused(self):
        head = self.model[-1]
        # the reference's fuse() swaps forward for fuseforward (yolo.py:705,708)
        return getattr(getattr(head, 'forward', None), '__name__', '') == 'fuseforward'


def accelerate(ref_model):
    """Return a drop-in module whose forward runs on the B200 engine using ``ref_model``'s weights."""
    return AcceleratedModel(ref_model)


class Ensemble(nn.ModuleList):
    """Drop-in for models/experimental.py:68-81: NMS ensemble = the members' predictions concatenated along dim 1.
    Every member decodes straight into its row range of one ``pred`` buffer (no ``torch.cat`` copy)."""

    def __init__(self, models=()):
        super().__init__(models)

    @torch.no_grad()
    def forward(self, x, augment=False):
        if not len(self):
            raise RuntimeError('Ensemble is empty')
        B, _, H, W = x.shape
        rows = [m.rows_per_image(H, W, augment) for m in self]
        no = self[0]._layers()[-1].no
        if any(m._layers()[-1].no != no for m in self):
            raise ValueError('Ensemble members must share nc (torch.cat along dim 1, experimental.py:80)')
        dev = x.device if x.device.type == 'cuda' else self[0]._device()
        pred = torch.empty((B, sum(rows), no), dtype=torch.float32, device=dev)
        off = 0
        for m, r in zip(self, rows):
            assert m.forward_into(x, pred, off, augment) == r
            off += r
        return pred, None

    @property
    def stride(self):
        return self[-1].stride

    @property
    def names(self):
        return self[-1].names
