#!/bin/bash
# the no-allocation entry point: alone in a fresh process, then after every other GPU test (the round-end order)
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests/test_zz_no_alloc.py 2>&1 | tail -4
timeout 1500 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider 2>&1 | tail -6
