"""Where the time between two dependent k_conv_tc launches goes: per-CTA globaltimer stamps (env YB_CONV_TS, read at
plan creation) of a producer layer A and a consumer layer B (B reads A's output), launched back to back on one stream
the way the engine does (programmatic dependent launch).  Usage:
  python scripts/conv_chain_timeline.py "CinA,CoutA,kA,sA,H,B" "CoutB,kB,sB" [...]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import _lib, ops  # noqa: E402

NAMES = ['entry', 'setup done', 'pdl_wait done', 'first TMA issued', 'first operands landed', 'first tile MMAs committed',
         'first accumulator ready', 'first TMA store issued', 'store smem read', 'exit']
L = _lib.lib()


def plan_for(x, co, k, s, ts):
    B, h, _, ci = x.shape
    os.environ['YB_CONV_TS'] = hex(ts.data_ptr())
    w = torch.randn((co, ci, k, k)) / (ci * k * k) ** 0.5
    out = torch.zeros((B, h // s, h // s, co), dtype=torch.bfloat16, device='cuda')
    wp = ops.pack_weight(w.cuda(), ci, torch.bfloat16)
    bp = torch.zeros(co, device='cuda')
    d = _lib.ConvDesc()
    d.x, d.y = _lib.view(x), _lib.view(out, 0, co)
    d.weight, d.bias = wp.data_ptr(), bp.data_ptr()
    d.cin_pad, d.cout_pad, d.k, d.stride, d.act, d.upsample, d.tile_n = ci, co, k, s, 1, 1, -1
    plan = C.c_void_p()
    _lib.check(L.yb_conv_plan_create(C.byref(d), C.byref(plan)))
    return plan, out, (wp, bp)


def col(t, i):
    v = t[:, i]
    return v[v > 0]


args = sys.argv[1:]
for a, b in zip(args[0::2], args[1::2]):
    ci, co, k, s, h, B = [int(t) for t in a.split(',')]
    co2, k2, s2 = [int(t) for t in b.split(',')]
    x = (torch.randn((B, h, h, ci), device='cuda') * 0.5).bfloat16()
    tsa = torch.zeros(16 * 1024, dtype=torch.int64, device='cuda')
    tsb = torch.zeros(16 * 1024, dtype=torch.int64, device='cuda')
    pa, ya, keep_a = plan_for(x, co, k, s, tsa)
    pb, yb, keep_b = plan_for(ya, co2, k2, s2, tsb)
    for it in range(5):
        tsa.zero_(), tsb.zero_()
        for _ in range(3):
            _lib.check(L.yb_conv_plan_launch(pa, _lib.stream_ptr()))
            _lib.check(L.yb_conv_plan_launch(pb, _lib.stream_ptr()))
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        _lib.check(L.yb_conv_plan_launch(pa, _lib.stream_ptr()))
        _lib.check(L.yb_conv_plan_launch(pb, _lib.stream_ptr()))
    e1.record()
    torch.cuda.synchronize()
    per_pair = 1e3 * e0.elapsed_time(e1) / 30
    ta, tb = tsa.view(-1, 16).cpu(), tsb.view(-1, 16).cpu()
    ta, tb = ta[ta[:, 0] > 0], tb[tb[:, 0] > 0]
    t0 = int(ta[:, 0].min())
    print(f'A = {a}, B = {b}: {per_pair:.2f} us per A+B pair; grid A {ta.shape[0]} CTAs, B {tb.shape[0]} CTAs; ns since A\'s first entry (min / median / max over CTAs)')
    for tag, t in (('A', ta), ('B', tb)):
        for i, n in enumerate(NAMES):
            v = col(t, i)
            if v.numel():
                v = v - t0
                print(f'   {tag} {n:28s} {int(v.min()):8d} {int(v.median()):8d} {int(v.max()):8d}')
    L.yb_conv_plan_destroy(pa), L.yb_conv_plan_destroy(pb)
