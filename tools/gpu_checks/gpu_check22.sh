#!/bin/bash
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_packed_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/bench_gemm.py 2048 2>&1 | tail -6
