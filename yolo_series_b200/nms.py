"""Drop-in ``non_max_suppression`` (reference utils/general.py:628-720) on the B200 NMS kernels.

Signature, defaults and return type are the reference's; ``max_det`` / ``max_nms`` (hard-coded
locals 300 / 30000 at general.py:641-642) are appended as trailing keywords.  Differences, all
documented in SURVEY.md A.6: no ``time_limit`` abort; distinct empty tensors instead of one shared
object; computation always in fp32 (the reference's fp16-on-CUDA class-offset overflow is not copied).
"""
import ctypes as C

import torch

from . import _lib

_ws_cache = {}


def _workspace(nbytes, device, slot=None):
    """Scratch buffer of the calling stream: two streams may run the NMS kernels at the same time (Model.detect_batches
    keeps two batches in flight), so the cache is keyed by (device, current stream)."""
    if slot is None:
        slot = torch.cuda.current_stream(device).cuda_stream
    key = (str(device), slot)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def _run_nms(pred, B, A, no, conf_thres, iou_thres, classes, agnostic, multi_label, max_nms, max_det, device,
             levels=None, cand_cap=0, return_packed=False, peer=None):
    """``peer``: a ``_lib.PeerGatherC`` table (dist.PeerGather.table) -- the fused path then also stores every kept row
    into the receive buffers of all GPUs of the box (yb_decode_nms_gather)."""
    L = _lib.lib()
    nbytes = L.yb_nms_workspace_bytes(B, A, no, int(bool(multi_label)), int(cand_cap))
    ws = _workspace(nbytes, device)
    out = torch.empty((B, max_det, 6), dtype=torch.float32, device=device)
    meta = torch.zeros(B + 1, dtype=torch.int32, device=device)   # counts[B] + overflow flag
    cls_arr, n_cls = None, 0
    if classes is not None:
        cl = [int(c) for c in (classes.tolist() if hasattr(classes, 'tolist') else classes)]
        n_cls = len(cl)
        cls_arr = (C.c_int * max(n_cls, 1))(*cl)
    ws_ptr = (ws.data_ptr() + 255) // 256 * 256
    ws_len = ws.numel() - (ws_ptr - ws.data_ptr())
    common = (C.c_float(conf_thres), C.c_double(iou_thres), cls_arr, n_cls, int(bool(agnostic)),
              int(bool(multi_label)), int(max_nms), int(max_det), int(cand_cap), C.c_void_p(ws_ptr), ws_len,
              C.c_void_p(out.data_ptr()), C.c_void_p(meta.data_ptr()), C.c_void_p(meta.data_ptr() + 4 * B))
    if levels is None and peer is not None:
        raise ValueError('the peer gather is fused into yb_decode_nms_gather (Model.detect), not the stand-alone yb_nms')
    with torch.cuda.device(device):          # the C side launches on the current device
        if levels is None:
            _lib.check(L.yb_nms(C.c_void_p(pred.data_ptr()), B, A, no, *common, _lib.stream_ptr()), 'yb_nms')
        elif peer is not None:
            lv, nl, na = levels
            _lib.check(L.yb_decode_nms_gather(lv, nl, na, no, *common, C.byref(peer), _lib.stream_ptr()),
                       'yb_decode_nms_gather')
        else:
            lv, nl, na = levels
            _lib.check(L.yb_decode_nms(lv, nl, na, no, *common, _lib.stream_ptr()), 'yb_decode_nms')
    if return_packed:
        return out, meta
    host = meta.cpu()   # the single device->host read of the call (B+1 ints)
    if int(host[B]) != 0:
        raise RuntimeError('yb_nms: candidate capacity exceeded (YB_ERR_OVERFLOW); raise cand_cap')
    counts = host[:B].tolist()
    return [out[i, :counts[i]] for i in range(B)]


def non_max_suppression(prediction, conf_thres=0.25, iou_thres=0.45, classes=None, agnostic=False,
                        multi_label=False, labels=(), max_det=300, max_nms=30000):
    """Runs NMS on inference results.  Returns list of detections, one (n,6) tensor per image
    [xyxy, conf, cls]."""
    if prediction.device.type != 'cuda':
        raise RuntimeError('yolo_series_b200.non_max_suppression needs a CUDA tensor (no CPU fallback)')
    in_dtype = prediction.dtype
    pred = prediction.float().contiguous()
    B, A, no = pred.shape
    if labels and len(labels):
        # a-priori label rows are appended after the image's own rows (general.py:656-662); rows with
        # obj == 0 are padding and never pass the obj > conf_thres filter.
        nl_max = max(len(l) for l in labels)
        if nl_max:
            extra = torch.zeros((B, nl_max, no), dtype=torch.float32, device=pred.device)
            for xi, l in enumerate(labels):
                if len(l):
                    l = torch.as_tensor(l, dtype=torch.float32, device=pred.device)
                    extra[xi, :len(l), :4] = l[:, 1:5]
                    extra[xi, :len(l), 4] = 1.0
                    extra[xi, torch.arange(len(l)), l[:, 0].long() + 5] = 1.0
            pred = torch.cat((pred, extra), 1).contiguous()
            A = pred.shape[1]
    out = _run_nms(pred, B, A, no, float(conf_thres), float(iou_thres), classes, agnostic, multi_label, max_nms,
                   max_det, pred.device)
    if in_dtype != torch.float32:
        out = [o.to(in_dtype) for o in out]
    return out
