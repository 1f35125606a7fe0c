#!/bin/bash
timeout 900 python -m pytest tests/test_export_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/bench_gemm.py 2>&1 | tail -8
