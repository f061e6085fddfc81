"""yb_nms_static: the no-allocation variant of yb_nms (scratch in a library-owned static __device__ array; no counterpart
in the reference).  Called straight through the C ABI (ctypes) with
caller-allocated inputs / outputs and compared with the existing entry point yb_nms -- bit-exact, the bar of every NMS
test here -- on two calls in a row with different shapes and modes (the array keeps the previous call's contents).
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle.synth import synth_pred

pytestmark = pytest.mark.gpu


def _static_nms(pred, conf, iou, multi_label, max_det, out, meta):
    """One call of yb_nms_static; `out` / `meta` are allocated by the caller beforehand (the entry point allocates nothing)."""
    from yolo_series_b200 import _lib
    L = _lib.lib()
    B, A, no = pred.shape
    rc = L.yb_nms_static(C.c_void_p(pred.data_ptr()), B, A, no, C.c_float(conf), C.c_double(iou), None, 0, 0,
                         int(multi_label), 30000, max_det, C.c_void_p(out.data_ptr()), C.c_void_p(meta.data_ptr()),
                         C.c_void_p(meta.data_ptr() + 4 * B), _lib.stream_ptr())
    return rc


def _rows(out, meta, B):
    cnt = meta[:B].tolist()
    return cnt, np.concatenate([out[i, :cnt[i]].cpu().numpy() for i in range(B)], 0)


def test_nms_static_equals_nms_on_two_calls_in_a_row():
    from yolo_series_b200 import _lib
    from yolo_series_b200.nms import _run_nms
    L = _lib.lib()
    dev = torch.device('cuda', torch.cuda.current_device())
    assert L.yb_nms_static_capacity() == 16 << 20
    cases = [   # (B, A, nc, seed, conf, iou, multi_label, max_det): the second call is smaller and in another mode
        (4, 20000, 80, 41, 0.001, 0.65, False, 300),
        (2, 1500, 80, 42, 0.05, 0.5, True, 100),
        (4, 20000, 80, 41, 0.001, 0.65, False, 300),      # and the first one again after the array has been reused
    ]
    preds = [synth_pred(B, A, nc, seed=seed, p_obj=0.3, clusters=(100, 200), extent=1280.0).to(dev)
             for B, A, nc, seed, *_ in cases]
    outs = [(torch.zeros((c[0], c[7], 6), dtype=torch.float32, device=dev),
             torch.zeros(c[0] + 1, dtype=torch.int32, device=dev)) for c in cases]
    torch.cuda.synchronize()
    for (B, A, nc, seed, conf, iou, ml, md), pred, (out, meta) in zip(cases, preds, outs):     # back to back, one stream
        assert _static_nms(pred, conf, iou, ml, md, out, meta) == 0, L.yb_last_error()
    torch.cuda.synchronize()
    total = 0
    for (B, A, nc, seed, conf, iou, ml, md), pred, (out, meta) in zip(cases, preds, outs):
        want_out, want_meta = _run_nms(pred, B, A, nc + 5, conf, iou, None, False, ml, 30000, md, dev, return_packed=True)
        assert int(meta[B]) == 0 and int(want_meta[B]) == 0
        cnt, rows = _rows(out, meta, B)
        wcnt, wrows = _rows(want_out, want_meta, B)
        assert cnt == wcnt
        assert np.array_equal(rows, wrows)
        total += sum(cnt)
    assert total > 0


def test_nms_static_rejects_sizes_beyond_the_static_array():
    from yolo_series_b200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda', torch.cuda.current_device())
    B, A, no = 64, 25200, 85                     # 64 * 25 200 * 18 B = 29 MB > 16 MiB
    pred = torch.zeros((1, 8, no), device=dev)   # never read: the size check comes before any launch
    out = torch.zeros((B, 10, 6), device=dev)
    meta = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    rc = L.yb_nms_static(C.c_void_p(pred.data_ptr()), B, A, no, C.c_float(0.25), C.c_double(0.45), None, 0, 0, 0, 30000,
                         10, C.c_void_p(out.data_ptr()), C.c_void_p(meta.data_ptr()), None, _lib.stream_ptr())
    assert rc == -3                              # YB_ERR_WORKSPACE
    assert b'static array holds' in L.yb_last_error()
    # multi_label blows the candidate capacity up by nc: a shape that fits in best-class mode is rejected there
    rc = L.yb_nms_static(C.c_void_p(pred.data_ptr()), 4, 20000, no, C.c_float(0.25), C.c_double(0.45), None, 0, 0, 1,
                         30000, 10, C.c_void_p(out.data_ptr()), C.c_void_p(meta.data_ptr()), None, _lib.stream_ptr())
    assert rc == -3
    torch.cuda.synchronize()
