#!/bin/bash
timeout 900 python -m pytest tests/test_fakequant_gpu.py tests/test_autoclip_gpu.py -m gpu -x -q 2>&1 | tail -3
echo "=== new"; timeout 300 python tools/bench_nf3.py 2>&1 | tail -5
echo "=== old (op-by-op)"; BDQ_NF3_THR=0 timeout 300 python tools/bench_nf3.py 2>&1 | tail -5
