"""BASELINE configs[4] alone (for ncu launch lists): python scripts/nms_stress_only.py"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

if __name__ == '__main__':
    torch.cuda.set_device(0)
    peaks, _ = bench.load_peaks()
    print(json.dumps(bench.nms_stress(torch.device('cuda', 0), peaks)))
