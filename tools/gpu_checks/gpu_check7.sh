#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > gpurun_out/pytest_gemm.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gemm.log
tail -30 gpurun_out/pytest_gemm.log
timeout 300 python tools/bench_gemm.py 2048 > gpurun_out/gemm_2048.log 2>&1; tail -6 gpurun_out/gemm_2048.log
