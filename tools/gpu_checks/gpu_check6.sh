#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_autoclip_gpu.py -m gpu -x -q -s > gpurun_out/pytest_ac.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_ac.log
tail -40 gpurun_out/pytest_ac.log
timeout 300 python tools/bench_autoclip.py int > gpurun_out/autoclip_int.log 2>&1; tail -5 gpurun_out/autoclip_int.log
timeout 300 python tools/bench_autoclip.py nf3 > gpurun_out/autoclip_nf3.log 2>&1; tail -5 gpurun_out/autoclip_nf3.log
