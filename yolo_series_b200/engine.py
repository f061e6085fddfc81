"""Shape-specialised instance of a lowered graph: HBM buffers, packed weights, C-side launch list.

Memory layout (DESIGN.md "data layout in HBM"): every root tensor is its own contiguous NHWC
buffer ``(B, H, W, C)`` in bf16 (trunk) or fp32 (fp32 reference mode; head logits are always fp32);
concat members are channel slices of their consumer's buffer.  Nothing is re-used between layers:
at yolov7/640/B=32 the whole activation set is 5.5 GB of the 180 GB HBM3e, so a liveness planner
would only add risk.  Weights are packed ``[Cout][kh*kw][Cin_pad]`` in the trunk dtype, bias fp32.
"""
import ctypes as C
import os

import torch

from . import _lib
from ._lib import OP_ADD, OP_CONV, OP_MAXPOOL2, OP_STEM, YB_BF16, YB_F32, YB_U8


def _round_up(x, m):
    return (x + m - 1) // m * m


class Engine:
    def __init__(self, lowered, B, H, W, device, dtype=torch.bfloat16, use_graph=True, tile_overrides=None,
                 img_dtype=torch.uint8, fuse_stem=True, packed=None):
        if device.type != 'cuda':
            raise RuntimeError('yolo_series_b200 runs on CUDA (sm_100a) only; there is no CPU fallback')
        if device.index is None:
            device = torch.device('cuda', torch.cuda.current_device())
        # every C entry point launches on the CURRENT device: pin it to the engine's device for all calls
        with torch.cuda.device(device):
            self._build(lowered, B, H, W, device, dtype, use_graph, tile_overrides, img_dtype, fuse_stem, packed)

    def _build(self, lowered, B, H, W, device, dtype, use_graph, tile_overrides, img_dtype, fuse_stem, packed):
        self.L, self.B, self.H, self.W = lowered, B, H, W
        self.device, self.dtype, self.use_graph = device, dtype, use_graph
        self.lib = _lib.lib()
        L = lowered
        cpad_in = 16 if dtype == torch.bfloat16 else 4
        # ---- buffers
        self.buffers = {}
        self.views = []

        def root_of(v):
            off = 0
            while L.vals[v].parent is not None:
                off += L.vals[v].offset
                v = L.vals[v].parent
            return v, off

        self.ops, self.pool_of = self._fuse_elementwise(L, dtype)
        used = set(self.pool_of.values())
        for op in self.ops:
            used.update(op.ins)
            used.update(op.outs)
        used.add(L.input_val)
        used.update(L.head_vals)
        used.update(L.aux_vals)
        for v in L.vals:
            if v.vid not in used:
                self.views.append(None)
                continue
            r, off = root_of(v.vid)
            rv = L.vals[r]
            if r not in self.buffers:
                h, w = H * rv.num // rv.den, W * rv.num // rv.den
                if r == L.input_val:
                    cb = _round_up(rv.C, cpad_in)
                else:
                    cb = _round_up(rv.C, 8)
                dt = torch.float32 if rv.f32 else dtype
                self.buffers[r] = torch.zeros((B, h, w, cb), dtype=dt, device=device)
            buf = self.buffers[r]
            cview = buf.shape[3] if v.vid == L.input_val else v.C
            self.views.append(_lib.view(buf, off, cview))
        # ---- fused staging + first conv (bf16 path): the stem kernel reads the raw NCHW image itself
        self.img_buf, stem_op = None, None
        if dtype == torch.bfloat16 and fuse_stem:
            readers = [op for op in self.ops if L.input_val in op.ins]
            if (len(readers) == 1 and readers[0].kind == OP_CONV and readers[0].conv.k == 3 and readers[0].upsample == 1
                    and readers[0].conv.c_in <= 16 and readers[0].conv.c_out % 16 == 0 and readers[0].conv.c_out <= 96
                    and not L.vals[readers[0].outs[0]].f32):
                stem_op = readers[0]
                self.img_buf = torch.zeros((B, L.in_ch, H, W), dtype=img_dtype, device=device)
                self.views.append(_lib.Tensor(self.img_buf.data_ptr(), _lib.torch_dtype_code(img_dtype), B, H, W,
                                              L.in_ch, L.in_ch))
        self.img_dtype = img_dtype
        # ---- weights
        self.weights, self.biases, self._keep = [], [], []
        ops_c = (_lib.Op * len(self.ops))()
        for i, op in enumerate(self.ops):
            o = ops_c[i]
            o.kind = op.kind
            o.in0 = op.ins[0]
            o.in1 = op.ins[1] if len(op.ins) > 1 else -1     # OP_ADD: second term; OP_CONV: residual (fused Shortcut)
            o.out0 = op.outs[0]
            o.out1 = op.outs[1] if len(op.outs) > 1 else -1
            o.out2 = op.outs[2] if len(op.outs) > 2 else -1
            o.k, o.stride, o.act, o.upsample, o.leaky_slope, o.weight = op.k, 1, 0, op.upsample, 0.0, -1
            o.pool_out = self.pool_of.get(id(op), -1)
            if op.kind == OP_CONV:
                fc = op.conv
                xin = self.views[op.ins[0]]
                cin_pad = max(_round_up(fc.c_in, 8 if dtype == torch.bfloat16 else 4), xin.C)
                if dtype == torch.bfloat16:
                    cin_pad = max(cin_pad, 16)
                # packed weights do not depend on the input shape: engines of one model share them (`packed` is the
                # model's cache, keyed by op name / padding / dtype / device)
                pk = (op.name, cin_pad, dtype, str(device))
                if packed is not None and pk in packed:
                    wd, bd = packed[pk]
                else:
                    w = torch.zeros((fc.c_out, fc.k * fc.k, cin_pad), dtype=torch.float32)
                    w[:, :, :fc.c_in] = fc.weight.permute(0, 2, 3, 1).reshape(fc.c_out, fc.k * fc.k, fc.c_in)
                    wd = w.to(device=device, dtype=dtype).contiguous()
                    bd = fc.bias.to(device=device, dtype=torch.float32).contiguous()
                    if packed is not None:
                        packed[pk] = (wd, bd)
                self.weights.append(wd)
                self.biases.append(bd)
                o.weight = len(self.weights) - 1
                o.k, o.stride, o.act, o.leaky_slope = fc.k, fc.s, fc.act, fc.slope
                o.cin_pad, o.cout_pad = cin_pad, fc.c_out
                if tile_overrides and op.name in tile_overrides:
                    ov = tuple(tile_overrides[op.name]) + (0,)
                    o.tile_n, o.stages, o.pair = ov[0], ov[1], ov[2]
                if op is stem_op:
                    o.kind, o.in0, o.tile_n = OP_STEM, len(self.views) - 1, int(L.reorg)
        tens_c = (_lib.Tensor * len(self.views))()
        for i, v in enumerate(self.views):
            if v is not None:
                tens_c[i] = v
        nw = len(self.weights)
        w_ptrs = (C.c_void_p * max(nw, 1))(*[w.data_ptr() for w in self.weights])
        b_ptrs = (C.c_void_p * max(nw, 1))(*[b.data_ptr() for b in self.biases])
        handle = C.c_void_p()
        _lib.check(self.lib.yb_engine_create(tens_c, len(self.views), ops_c, len(self.ops), w_ptrs, b_ptrs, nw,
                                             C.byref(handle)), 'yb_engine_create')
        self.handle = handle
        self.n_ops = len(self.ops)
        self.head = [self.buffers[v] for v in L.head_vals]        # fp32 NHWC logits per level
        self.head_views = [self.views[v] for v in L.head_vals]
        self.input_view = self.views[L.input_val]
        self.conv_flops = self._count_flops()

    @staticmethod
    def _fuse_elementwise(L, dtype):
        """bf16 path: fold the element-wise neighbours of 1x1 convs into the conv kernel's epilogue.  Returns
        (ops, pool_of): the op list to run (copies where an op was changed -- the lowered graph is shared between
        engines) and {id(conv op): pooled value id}.

        * Shortcut (OP_ADD, models/common.py:80-86): ``a + b`` where the later-computed term is the output of a 1x1 conv
          that nobody else reads -> that conv adds the other term as a residual after its activation and writes the
          sum (yb_conv_desc.res); the add kernel and one round trip of the tensor disappear.
        * MP (OP_MAXPOOL2, common.py:30-36) whose input comes from a 1x1 conv -> the conv also writes the 2x2 max-pool
          of every tile it has staged (yb_conv_desc.pool); bit-identical to the separate pool kernel.
        """
        import dataclasses
        ops = list(L.ops)
        pool_of = {}
        if dtype != torch.bfloat16 or os.environ.get('YB_NO_EPILOGUE_FUSION') is not None:
            return ops, pool_of

        def fusable_conv(op):
            return (op is not None and op.kind == OP_CONV and op.conv.k == 1 and op.upsample == 1 and len(op.ins) == 1
                    and L.input_val not in op.ins and not L.vals[op.outs[0]].f32)
        if os.environ.get('YB_NO_RES_FUSION') is None:
            pos = {id(op): i for i, op in enumerate(ops)}
            producer = {v: op for op in ops for v in op.outs}
            readers = {}                     # live readers per value (Val.users still counts ops dropped as dead code)
            for op in ops:
                for v in op.ins:
                    readers[v] = readers.get(v, 0) + 1
            for v in list(L.head_vals) + list(L.aux_vals):
                readers[v] = readers.get(v, 0) + 1
            for add in [o for o in ops if o.kind == OP_ADD]:
                a, b = add.ins[0], add.ins[1]
                pa, pb = producer.get(a), producer.get(b)
                if pa is None or pb is None:
                    continue
                # the term computed LAST carries the add; the other one is ready by then
                (late, lv), (early_v) = ((pa, a), b) if pos[id(pa)] > pos[id(pb)] else ((pb, b), a)
                v = L.vals[lv]
                if not fusable_conv(late) or readers.get(lv, 0) != 1 or v.parent is not None or v.is_concat:
                    continue
                fused = dataclasses.replace(late, ins=[late.ins[0], early_v], outs=list(add.outs), name=late.name + '+add')
                i = pos[id(late)]
                ops[i] = fused
                ops.remove(add)
                pos = {id(op): k for k, op in enumerate(ops)}
                producer = {vv: op for op in ops for vv in op.outs}
        if os.environ.get('YB_NO_POOL_FUSION') is None:
            producer = {v: op for op in ops for v in op.outs}
            for mp in [o for o in ops if o.kind == OP_MAXPOOL2]:
                src = producer.get(mp.ins[0])
                if fusable_conv(src) or (src is not None and src.kind == OP_CONV and src.conv.k == 1 and src.upsample == 1
                                         and len(src.ins) == 2 and not L.vals[src.outs[0]].f32):
                    if id(src) not in pool_of:
                        pool_of[id(src)] = mp.outs[0]
                        ops.remove(mp)
        return ops, pool_of

    def _count_flops(self):
        total = 0
        for op in self.L.ops:
            if op.kind == OP_CONV:
                fc = op.conv
                v = self.L.vals[op.ins[0]]
                ho = (self.H * v.num // v.den) // fc.s
                wo = (self.W * v.num // v.den) // fc.s
                total += 2 * fc.c_in * fc.c_out * fc.k * fc.k * ho * wo
        return total  # per image

    def stage(self, img):
        """img: (B, C, H, W) uint8 or float32 CUDA tensor -> staged NHWC input buffer."""
        if img.dtype not in (torch.uint8, torch.float32):
            img = img.float()
        if self.img_buf is not None:
            if img.dtype != self.img_dtype or tuple(img.shape) != tuple(self.img_buf.shape):
                raise ValueError(f'engine built for {tuple(self.img_buf.shape)} {self.img_dtype}, got '
                                 f'{tuple(img.shape)} {img.dtype}')
            with torch.cuda.device(self.device):
                self.img_buf.copy_(img, non_blocking=True)     # D2D, or H2D straight from (pinned) host memory
            return
        if img.device != self.device:
            img = img.to(self.device, non_blocking=True)
        img = img.contiguous()
        code = YB_U8 if img.dtype == torch.uint8 else YB_F32
        b, c, h, w = img.shape
        if (b, h, w) != (self.B, self.H, self.W):
            raise ValueError(f'engine built for {(self.B, self.H, self.W)}, got {(b, h, w)}')
        with torch.cuda.device(self.device):
            _lib.check(self.lib.yb_stage_input(C.c_void_p(img.data_ptr()), code, b, c, h, w, int(self.L.reorg),
                                               C.byref(self.input_view), _lib.stream_ptr()), 'yb_stage_input')

    def run(self):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.yb_engine_forward(self.handle, int(self.use_graph), _lib.stream_ptr()),
                       'yb_engine_forward')

    def profile(self):
        """Per-op device time (ms) of one eager pass, with op descriptions: [(name, kind, ms, flops)]."""
        ms = (C.c_float * self.n_ops)()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.yb_engine_profile(self.handle, ms, _lib.stream_ptr()), 'yb_engine_profile')
        rows = []
        for i, op in enumerate(self.ops):
            fl = 0
            if op.kind == OP_CONV:
                fc = op.conv
                v = self.L.vals[op.ins[0]]
                ho = (self.H * v.num // v.den) // fc.s
                wo = (self.W * v.num // v.den) // fc.s
                fl = 2 * fc.c_in * fc.c_out * fc.k * fc.k * ho * wo * self.B
            rows.append((op.name, op.kind, float(ms[i]), fl))
        return rows

    def op_kernel(self, i):
        """Name of the CUDA kernel op ``i`` launches (yb_engine_op_kernel)."""
        n = self.lib.yb_engine_op_kernel(self.handle, i)
        return n.decode() if n else '?'

    def op_bytes(self, i):
        """Algorithmic HBM bytes of op ``i``: every input and output element once, weights once."""
        op = self.ops[i]

        def vb(v, c=None):
            return v.N * v.H * v.W * (v.C if c is None else c) * {YB_F32: 4, YB_BF16: 2, YB_U8: 1}[v.dtype]
        total = sum(vb(self.views[o]) for o in op.outs)
        if id(op) in self.pool_of:
            total += vb(self.views[self.pool_of[id(op)]])
        if op.kind == OP_CONV and self.img_buf is not None and self.L.input_val in op.ins:     # fused stem reads the raw image
            total += self.img_buf.numel() * self.img_buf.element_size()
        else:
            total += sum(vb(self.views[v]) for v in op.ins)
        if op.kind == OP_CONV:
            fc = op.conv
            total += fc.c_out * fc.k * fc.k * fc.c_in * (2 if self.dtype == torch.bfloat16 else 4) + fc.c_out * 4
        return total

    def profile_table(self, reps=3):
        """[{op, kernel, ms, flops, bytes}] -- per-op device time of an eager pass (median of ``reps``)."""
        runs = [self.profile() for _ in range(reps)]
        rows = []
        for i, r in enumerate(runs[0]):
            ms = sorted(run[i][2] for run in runs)[reps // 2]
            rows.append({'op': r[0], 'kernel': self.op_kernel(i), 'ms': ms, 'flops': r[3], 'bytes': self.op_bytes(i)})
        return rows

    def num_launches(self):
        return self.lib.yb_engine_num_launches(self.handle)

    def destroy(self):
        """Free the C-side plans / graph now (the HBM buffers go when the last Python reference does)."""
        h = getattr(self, 'handle', None)
        if h:
            try:
                with torch.cuda.device(self.device):
                    torch.cuda.synchronize(self.device)       # a replay may still be in flight
                    self.lib.yb_engine_destroy(h)
            except Exception:
                pass
            self.handle = None
        self.buffers, self.views, self.head = {}, [], []

    def __del__(self):
        self.destroy()
