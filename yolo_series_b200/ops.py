"""Thin per-operator Python wrappers over the C ABI (NHWC torch tensors in, NHWC out).
Used by the parity tests and micro-benchmarks; Model.forward goes through engine.py instead."""
import ctypes as C

import torch

from . import _lib
from ._lib import ACT_NONE


def pack_weight(w, cin_pad, dtype):
    """(Cout, Cin, k, k) fp32 -> [Cout][k*k][cin_pad] in ``dtype`` (zero padded)."""
    co, ci, k, _ = w.shape
    out = torch.zeros((co, k * k, cin_pad), dtype=torch.float32, device=w.device)
    out[:, :, :ci] = w.permute(0, 2, 3, 1).reshape(co, k * k, ci)
    return out.to(dtype).contiguous()


def conv2d(x, w, b, stride=1, act=ACT_NONE, slope=0.0, upsample=1, out=None, out_c_off=0, out_dtype=None,
           tile_n=0, stages=0, pair=0, pool_out=None, res=None):
    """x: (N,H,W,Cx) NHWC cuda tensor (bf16 -> tcgen05 path, fp32 -> reference-mode path);
    w: (Cout, Cin, k, k) fp32 with Cin <= Cx; b: (Cout,) fp32."""
    L = _lib.lib()
    n, h, wd, cx = x.shape
    co, ci, k, _ = w.shape
    pad = k // 2
    ho, wo = (h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1
    cin_pad = max(cx, 16 if x.dtype == torch.bfloat16 else 4)
    wp = pack_weight(w.to(x.device), cin_pad, x.dtype)
    bp = b.to(x.device, torch.float32).contiguous()
    if out is None:
        od = out_dtype or x.dtype
        cb = (co + 7) // 8 * 8
        out = torch.zeros((n, ho * upsample, wo * upsample, cb), dtype=od, device=x.device)
    d = _lib.ConvDesc()
    d.x = _lib.view(x)
    d.y = _lib.view(out, out_c_off, co)
    d.weight, d.bias = wp.data_ptr(), bp.data_ptr()
    d.cin_pad, d.cout_pad, d.k, d.stride = cin_pad, co, k, stride
    d.act, d.leaky_slope, d.upsample = act, slope, upsample
    d.tile_n, d.stages, d.pair = tile_n, stages, pair
    if res is not None:           # y = act(conv) + res (Shortcut folded into the conv)
        d.res = _lib.view(res, 0, co)
    if pool_out is not None:      # fused MaxPool2d(2, 2) of the output (second tensor written by the same kernel)
        d.pool = _lib.view(pool_out, 0, co)
    _lib.check(L.yb_conv2d(C.byref(d), _lib.stream_ptr()), 'yb_conv2d')
    return out


def _unary(fn_name, x, out, *extra):
    L = _lib.lib()
    vx, vy = _lib.view(x), _lib.view(out)
    _lib.check(getattr(L, fn_name)(C.byref(vx), C.byref(vy), *extra, _lib.stream_ptr()), fn_name)
    return out


def maxpool2x2(x):
    n, h, w, c = x.shape
    return _unary('yb_maxpool2x2', x, torch.empty((n, h // 2, w // 2, c), dtype=x.dtype, device=x.device))


def maxpool_s1(x, k):
    return _unary('yb_maxpool_s1', x, torch.empty_like(x), k)


def upsample2x(x):
    n, h, w, c = x.shape
    return _unary('yb_upsample2x', x, torch.empty((n, 2 * h, 2 * w, c), dtype=x.dtype, device=x.device))


def spp3(x):
    L = _lib.lib()
    n, h, w, c = x.shape
    out = torch.empty((n, h, w, 3 * c), dtype=x.dtype, device=x.device)
    vs = [_lib.view(out, i * c, c) for i in range(3)]
    vx = _lib.view(x)
    _lib.check(L.yb_spp3(C.byref(vx), C.byref(vs[0]), C.byref(vs[1]), C.byref(vs[2]), _lib.stream_ptr()), 'yb_spp3')
    return out


def add(a, b):
    L = _lib.lib()
    out = torch.empty_like(a)
    va, vb, vy = _lib.view(a), _lib.view(b), _lib.view(out)
    _lib.check(L.yb_add(C.byref(va), C.byref(vb), C.byref(vy), _lib.stream_ptr()), 'yb_add')
    return out


def copy_into(x, out, c_off):
    L = _lib.lib()
    vx, vy = _lib.view(x), _lib.view(out, c_off, x.shape[3])
    _lib.check(L.yb_copy(C.byref(vx), C.byref(vy), _lib.stream_ptr()), 'yb_copy')
    return out


def stage_input(img, cpad, dtype, reorg=False):
    L = _lib.lib()
    b, c, h, w = img.shape
    oh, ow = (h // 2, w // 2) if reorg else (h, w)
    out = torch.empty((b, oh, ow, cpad), dtype=dtype, device=img.device)
    vy = _lib.view(out)
    code = _lib.YB_U8 if img.dtype == torch.uint8 else _lib.YB_F32
    _lib.check(L.yb_stage_input(C.c_void_p(img.data_ptr()), code, b, c, h, w, int(reorg), C.byref(vy),
                                _lib.stream_ptr()), 'yb_stage_input')
    return out


def decode(logits, strides, anchor_grid, na, no, want_raw=True):
    """logits: list of NHWC fp32/bf16 (B,ny,nx,>=na*no); anchor_grid: (nl, na, 2) pixels."""
    L = _lib.lib()
    nl = len(logits)
    lv = (_lib.HeadLevel * nl)()
    B = logits[0].shape[0]
    A = 0
    for i, t in enumerate(logits):
        lv[i].logits = _lib.view(t)
        lv[i].stride = float(strides[i])
        for a in range(na):
            lv[i].anchor_wh[2 * a] = float(anchor_grid[i][a][0])
            lv[i].anchor_wh[2 * a + 1] = float(anchor_grid[i][a][1])
        A += na * t.shape[1] * t.shape[2]
    dev = logits[0].device
    pred = torch.empty((B, A, no), dtype=torch.float32, device=dev)
    raws = [torch.empty((B, na, t.shape[1], t.shape[2], no), dtype=torch.float32, device=dev) for t in logits]
    ptrs = (C.c_void_p * nl)(*[r.data_ptr() for r in raws]) if want_raw else None
    _lib.check(L.yb_decode(lv, nl, na, no, C.c_void_p(pred.data_ptr()), ptrs, _lib.stream_ptr()), 'yb_decode')
    return pred, raws, (lv, nl, na)


def stem_conv(img, w, b, stride=1, act=ACT_NONE, slope=0.0, reorg=False):
    """img: (B,C,H,W) uint8|fp32 cuda; w: (Cout, Cin', 3, 3) fp32 with Cin' = C (or 4C with reorg)."""
    L = _lib.lib()
    B, C_, H, W = img.shape
    co, ci = w.shape[0], w.shape[1]
    hi, wi = (H // 2, W // 2) if reorg else (H, W)
    ho, wo = (hi - 1) // stride + 1, (wi - 1) // stride + 1
    wp = pack_weight(w.to(img.device), 16, torch.bfloat16)
    bp = b.to(img.device, torch.float32).contiguous()
    out = torch.zeros((B, ho, wo, co), dtype=torch.bfloat16, device=img.device)
    vy = _lib.view(out)
    code = _lib.YB_U8 if img.dtype == torch.uint8 else _lib.YB_F32
    _lib.check(L.yb_stem_conv(C.c_void_p(img.data_ptr()), code, B, C_, H, W, int(reorg), C.c_void_p(wp.data_ptr()),
                              C.c_void_p(bp.data_ptr()), 16, stride, act, slope, C.byref(vy), _lib.stream_ptr()),
               'yb_stem_conv')
    return out
