#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fakequant_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/microbench.py > gpurun_out/microbench.log 2>&1; tail -8 gpurun_out/microbench.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json | cut -c1-900
