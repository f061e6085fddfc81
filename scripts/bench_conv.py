"""Micro-benchmark of single conv shapes through yb_conv2d-equivalent engine plans.
Usage: python scripts/bench_conv.py "Cin,Cout,k,s,H[,tile_n[,stages[,pair]]]" ... (B=32 fixed unless YB_B set)"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import _lib, ops  # noqa: E402

B = int(os.environ.get('YB_B', 32))
REPS = int(os.environ.get('YB_REPS', 20))
L = _lib.lib()
for spec in sys.argv[1:]:
    v = [int(t) for t in spec.split(',')]
    ci, co, k, s, h = v[:5]
    tile_n = v[5] if len(v) > 5 else 0
    stages = v[6] if len(v) > 6 else 0
    pair = v[7] if len(v) > 7 else 0
    cx = max(ci, 16)
    x = (torch.randn((B, h, h, cx), device='cuda') * 0.5).bfloat16()
    w = torch.randn((co, ci, k, k)) / (ci * k * k) ** 0.5
    b = torch.zeros(co)
    ho = h // s
    out = torch.zeros((B, ho, ho, (co + 7) // 8 * 8), dtype=torch.bfloat16, device='cuda')
    wp = ops.pack_weight(w.cuda(), cx, torch.bfloat16)
    bp = b.cuda()
    d = _lib.ConvDesc()
    d.x, d.y = _lib.view(x), _lib.view(out, 0, co)
    d.weight, d.bias = wp.data_ptr(), bp.data_ptr()
    d.cin_pad, d.cout_pad, d.k, d.stride, d.act, d.upsample = cx, co, k, s, 1, 1
    d.tile_n, d.stages, d.pair = tile_n, stages, pair
    plan = C.c_void_p()
    _lib.check(L.yb_conv_plan_create(C.byref(d), C.byref(plan)))
    for _ in range(3):
        _lib.check(L.yb_conv_plan_launch(plan, _lib.stream_ptr()))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(REPS):
        _lib.check(L.yb_conv_plan_launch(plan, _lib.stream_ptr()))
    e1.record()
    torch.cuda.synchronize()
    L.yb_conv_plan_destroy(plan)
    ms = e0.elapsed_time(e1) / REPS
    fl = 2 * ci * co * k * k * ho * ho * B
    by = B * (h * h * cx + ho * ho * co) * 2
    print(f'{spec}: {ms:.4f} ms  {fl / ms / 1e9:.1f} TFLOP/s  {by / ms / 1e6:.0f} GB/s algo', flush=True)
