#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_packed_gpu.py -m gpu -x -q > gpurun_out/pytest_packed.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_packed.log
tail -15 gpurun_out/pytest_packed.log
timeout 600 python tools/bench_gemv.py 1 > gpurun_out/gemv_b1.log 2>&1; cat gpurun_out/gemv_b1.log
