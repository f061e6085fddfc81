"""Timeline of one k_conv_tc launch (CTA 0) from in-kernel globaltimer stamps (env YB_CONV_TS = address of a device
buffer, read by conv_tc_plan_create).  Usage: python scripts/conv_timeline.py "Cin,Cout,k,s,H,B" ..."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from yolo_series_b200 import _lib, ops  # noqa: E402

NAMES = ['entry', 'setup done', 'pdl_wait done', 'first TMA issued', 'first operands landed', 'last MMA committed',
         'accumulator ready', 'TMA store issued', 'store smem read', 'exit']
L = _lib.lib()
ts = torch.zeros(16 * 1024, dtype=torch.int64, device='cuda')      # 16 stamps per CTA; CTA 0 is printed
os.environ['YB_CONV_TS'] = hex(ts.data_ptr())
for spec in sys.argv[1:]:
    ci, co, k, s, h, B = [int(t) for t in spec.split(',')]
    x = (torch.randn((B, h, h, ci), device='cuda') * 0.5).bfloat16()
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
    rows = []
    for it in range(6):
        for _ in range(4):                         # back to back: the measured launch has a predecessor in flight
            _lib.check(L.yb_conv_plan_launch(plan, _lib.stream_ptr()))
        torch.cuda.synchronize()
        rows.append(ts[:10].cpu().tolist())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        _lib.check(L.yb_conv_plan_launch(plan, _lib.stream_ptr()))
    e1.record()
    torch.cuda.synchronize()
    L.yb_conv_plan_destroy(plan)
    r = rows[-1]
    print(f'{spec}: {1e3 * e0.elapsed_time(e1) / 50:.2f} us per launch (50 back-to-back); last launch, ns since entry:')
    for n, v in zip(NAMES, r):
        print(f'   {n:24s} {v - r[0]:7d}')
