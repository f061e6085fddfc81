"""ctypes binding of libyolo_b200.so (C ABI declared in include/yolo_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C yolo_series_b200/csrc``.
There is NO fallback: if the shared object is missing, loading raises, and every compute entry
point raises ``RuntimeError`` with ``yb_last_error()`` on a non-zero status.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libyolo_b200.so')

YB_F32, YB_BF16, YB_U8 = 0, 1, 2
ACT_NONE, ACT_SILU, ACT_LEAKY = 0, 1, 2
OP_CONV, OP_MAXPOOL2, OP_POOL_S1, OP_SPP3, OP_UPSAMPLE2, OP_COPY, OP_ADD, OP_STEM = 1, 2, 3, 4, 5, 6, 7, 8
MAX_LEVELS, MAX_ANCHORS = 8, 8

EXPORTS = [
    'yb_last_error', 'yb_abi_version', 'yb_conv2d', 'yb_conv_plan_create', 'yb_conv_plan_launch',
    'yb_conv_plan_destroy', 'yb_maxpool2x2', 'yb_maxpool_s1', 'yb_spp3',
    'yb_upsample2x', 'yb_copy', 'yb_add', 'yb_stage_input', 'yb_stem_conv', 'yb_decode', 'yb_nms_workspace_bytes',
    'yb_nms', 'yb_nms_static', 'yb_nms_static_capacity', 'yb_decode_nms', 'yb_engine_create', 'yb_engine_forward', 'yb_engine_num_launches', 'yb_engine_profile',
    'yb_engine_op_kernel', 'yb_engine_destroy', 'yb_letterbox', 'yb_scale_coords', 'yb_scale_img', 'yb_decode_tta', 'yb_box_iou',
    'yb_match_detections', 'yb_decode_nms_gather', 'yb_peer_wait', 'yb_peer_alloc', 'yb_peer_open', 'yb_peer_close',
    'yb_peer_free',
]


class Tensor(C.Structure):
    _fields_ = [('data', C.c_void_p), ('dtype', C.c_int32), ('N', C.c_int32), ('H', C.c_int32),
                ('W', C.c_int32), ('C', C.c_int32), ('c_stride', C.c_int32)]


class ConvDesc(C.Structure):
    _fields_ = [('x', Tensor), ('y', Tensor), ('weight', C.c_void_p), ('bias', C.c_void_p),
                ('cin_pad', C.c_int32), ('cout_pad', C.c_int32), ('k', C.c_int32), ('stride', C.c_int32),
                ('act', C.c_int32), ('leaky_slope', C.c_float), ('upsample', C.c_int32), ('res', Tensor),
                ('tile_n', C.c_int32), ('stages', C.c_int32), ('pair', C.c_int32), ('pool', Tensor)]


class HeadLevel(C.Structure):
    _fields_ = [('logits', Tensor), ('stride', C.c_float), ('anchor_wh', C.c_float * (2 * MAX_ANCHORS))]


class Op(C.Structure):
    _fields_ = [('kind', C.c_int32), ('in0', C.c_int32), ('in1', C.c_int32), ('out0', C.c_int32),
                ('out1', C.c_int32), ('out2', C.c_int32), ('k', C.c_int32), ('stride', C.c_int32),
                ('act', C.c_int32), ('upsample', C.c_int32), ('leaky_slope', C.c_float), ('weight', C.c_int32),
                ('cin_pad', C.c_int32), ('cout_pad', C.c_int32), ('tile_n', C.c_int32), ('stages', C.c_int32),
                ('pair', C.c_int32), ('pool_out', C.c_int32)]


class LetterboxSrc(C.Structure):
    _fields_ = [('src', C.c_void_p), ('src_pitch', C.c_int64), ('src_h', C.c_int32), ('src_w', C.c_int32),
                ('new_w', C.c_int32), ('new_h', C.c_int32), ('top', C.c_int32), ('left', C.c_int32)]


class TtaPass(C.Structure):
    _fields_ = [('rows_total', C.c_int32), ('row_off', C.c_int32), ('scale', C.c_float), ('flip', C.c_int32),
                ('img_h', C.c_float), ('img_w', C.c_float)]


MAX_PEERS = 16


class PeerGatherC(C.Structure):
    _fields_ = [('n_peers', C.c_int32), ('rank', C.c_int32), ('out', C.c_void_p * MAX_PEERS),
                ('counts', C.c_void_p * MAX_PEERS), ('flags', C.c_void_p * MAX_PEERS), ('step', C.c_uint32),
                ('done_counter', C.c_void_p)]


LAYOUT_HWC, LAYOUT_CHW = 0, 1

_lib = None


def lib():
    """Load (once) and return the ctypes handle.  Raises if the extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f'{LIB_PATH} is missing: build it with `python -c "import __graft_entry__ as g; '
                           f'g.build()"` or `make -C yolo_series_b200/csrc`. There is no CPU fallback.')
    L = C.CDLL(LIB_PATH)
    L.yb_last_error.restype = C.c_char_p
    L.yb_nms_workspace_bytes.restype = C.c_size_t
    L.yb_nms_workspace_bytes.argtypes = [C.c_int] * 5
    L.yb_engine_destroy.restype = None
    L.yb_engine_destroy.argtypes = [C.c_void_p]
    L.yb_engine_forward.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.yb_engine_num_launches.argtypes = [C.c_void_p]
    L.yb_engine_op_kernel.argtypes = [C.c_void_p, C.c_int]
    L.yb_engine_op_kernel.restype = C.c_char_p
    L.yb_engine_profile.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_void_p]
    L.yb_conv2d.argtypes = [C.POINTER(ConvDesc), C.c_void_p]
    L.yb_conv_plan_create.argtypes = [C.POINTER(ConvDesc), C.POINTER(C.c_void_p)]
    L.yb_conv_plan_launch.argtypes = [C.c_void_p, C.c_void_p]
    L.yb_conv_plan_destroy.argtypes = [C.c_void_p]
    L.yb_conv_plan_destroy.restype = None
    for name in ('yb_maxpool2x2', 'yb_upsample2x', 'yb_copy'):
        getattr(L, name).argtypes = [C.POINTER(Tensor), C.POINTER(Tensor), C.c_void_p]
    L.yb_maxpool_s1.argtypes = [C.POINTER(Tensor), C.POINTER(Tensor), C.c_int, C.c_void_p]
    L.yb_spp3.argtypes = [C.POINTER(Tensor)] * 4 + [C.c_void_p]
    L.yb_add.argtypes = [C.POINTER(Tensor)] * 3 + [C.c_void_p]
    L.yb_stage_input.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.POINTER(Tensor), C.c_void_p]
    L.yb_stem_conv.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                               C.c_int, C.c_int, C.c_int, C.c_float, C.POINTER(Tensor), C.c_void_p]
    L.yb_decode.argtypes = [C.POINTER(HeadLevel), C.c_int, C.c_int, C.c_int, C.c_void_p,
                            C.POINTER(C.c_void_p), C.c_void_p]
    L.yb_nms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_double, C.POINTER(C.c_int), C.c_int,
                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                         C.c_void_p, C.c_void_p, C.c_void_p]
    L.yb_nms_static_capacity.restype = C.c_size_t
    L.yb_nms_static_capacity.argtypes = []
    L.yb_nms_static.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_double, C.POINTER(C.c_int), C.c_int,
                                C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.yb_decode_nms.argtypes = [C.POINTER(HeadLevel), C.c_int, C.c_int, C.c_int, C.c_float, C.c_double,
                                C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.yb_engine_create.argtypes = [C.POINTER(Tensor), C.c_int, C.POINTER(Op), C.c_int, C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p)]
    L.yb_letterbox.argtypes = [C.POINTER(LetterboxSrc), C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.POINTER(C.c_uint8), C.c_void_p]
    L.yb_scale_coords.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.yb_scale_img.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                               C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p]
    L.yb_decode_tta.argtypes = [C.POINTER(HeadLevel), C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(TtaPass),
                                C.c_void_p]
    L.yb_box_iou.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.yb_match_detections.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.yb_decode_nms_gather.argtypes = [C.POINTER(HeadLevel), C.c_int, C.c_int, C.c_int, C.c_float, C.c_double,
                                       C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.POINTER(PeerGatherC), C.c_void_p]
    L.yb_peer_wait.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_void_p]
    L.yb_peer_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]
    L.yb_peer_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.yb_peer_close.argtypes = [C.c_void_p]
    L.yb_peer_free.argtypes = [C.c_void_p]
    if L.yb_abi_version() != 4:
        raise RuntimeError('libyolo_b200.so ABI version mismatch; rebuild')
    _lib = L
    return L


def check(rc, what=''):
    if rc != 0:
        msg = lib().yb_last_error().decode('utf-8', 'replace')
        raise RuntimeError(f'{what or "libyolo_b200"} failed (status {rc}): {msg}')


def torch_dtype_code(t):
    import torch
    return {torch.float32: YB_F32, torch.bfloat16: YB_BF16, torch.uint8: YB_U8}[t]


def view(t, c_off=0, C_=None):
    """Describe a contiguous NHWC torch tensor (N,H,W,Cbuf) or a channel slice of it."""
    import torch
    if t.device.type != 'cuda' or t.device.index != torch.cuda.current_device():
        raise RuntimeError(f'tensor on {t.device} but the current CUDA device is cuda:{torch.cuda.current_device()}: '
                           f'libyolo_b200 launches on the current device -- wrap the call in torch.cuda.device(...)')
    n, h, w, cb = t.shape
    c = cb - c_off if C_ is None else C_
    return Tensor(t.data_ptr() + c_off * t.element_size(), torch_dtype_code(t.dtype), n, h, w, c, cb)


def stream_ptr(device=None):
    """Current CUDA stream of ``device`` (default: the current device).  Every C entry point launches on the CURRENT
    device, so callers that hold tensors of a specific device wrap the call in ``torch.cuda.device(dev)`` (Engine,
    nms._run_nms, preprocess, evaluate do) -- a model on cuda:1 must not launch on cuda:0."""
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
