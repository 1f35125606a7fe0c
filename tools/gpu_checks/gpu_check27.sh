#!/bin/bash
timeout 900 python -m pytest tests/test_fakequant_gpu.py -m gpu -x -q 2>&1 | tail -8
echo "=== new"; timeout 300 python tools/bench_nf3.py 2>&1 | tail -5
