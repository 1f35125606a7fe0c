#!/bin/bash
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_export_gpu.py -m gpu -x -q 2>&1 | tail -4
for f in tools/bench_gemm*.py; do timeout 600 python $f 2>&1 | tail -12; done
