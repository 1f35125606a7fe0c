#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_packed_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_gemv.py 1 > gpurun_out/gemv_b1.log 2>&1; tail -15 gpurun_out/gemv_b1.log
timeout 300 python tools/bench_gemv.py 8 > gpurun_out/gemv_b8.log 2>&1; tail -6 gpurun_out/gemv_b8.log
