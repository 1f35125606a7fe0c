#!/bin/bash
timeout 900 python -m pytest tests/test_packed_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/bench_gemv.py 1 2>&1 | grep "fused\|best"
echo "=== layer tma"; timeout 300 python tools/gemv_timeline.py layer 17 2>&1 | tail -5
