#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fakequant_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python tools/microbench.py > gpurun_out/microbench.log 2>&1; tail -20 gpurun_out/microbench.log
