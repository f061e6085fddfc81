"""yolo_series_b200 -- B200-native (sm_100a) YOLOv7 inference hot path behind the reference's
own Python API (qbxlvnf11/yolo-series): ``Model(cfg).forward(x)`` and ``non_max_suppression(...)``.

The importable package directory is ``yolo_series_b200`` (a hyphen, as in "yolo-series_b200", is
not a legal Python identifier).  All compute lives in ``libyolo_b200.so`` (csrc/, C ABI in
include/yolo_b200.h); importing the package does not load it, calling any operator does and raises
if it is missing -- there is no CPU fallback.
"""
from .model import Ensemble, Model, accelerate, parse_model  # noqa: F401
from .nms import non_max_suppression                        # noqa: F401
from .zoo import get_cfg                                    # noqa: F401

__all__ = ['Model', 'Ensemble', 'accelerate', 'parse_model', 'non_max_suppression', 'get_cfg']
