#!/bin/bash
timeout 900 python -m pytest tests/test_packed_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/gemv_variants.py 1 17 2>&1 | tail -3
echo "=== big"; timeout 300 python tools/gemv_big.py 1 2>&1 | tail -5
