#!/bin/bash
# Full GPU pass: parity tests, smoke, bench both arms, GEMV bench, ncu launch list of bench + full capture of the GEMV.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python tools/bench_gemv.py 1 > gpurun_out/gemv_b1.log 2>&1; grep "fused\|best" gpurun_out/gemv_b1.log | grep -v "ring\|1cta"
timeout 300 python tools/bench_gemv.py 4 > gpurun_out/gemv_b4.log 2>&1; tail -3 gpurun_out/gemv_b4.log
python tools/prof_gemv.py > gpurun_out/plain_gemv.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemv_w2s -s 4 -c 3 -f -o gpurun_out/prof_gemv_s2 python tools/prof_gemv.py > gpurun_out/ncu_gemv.log 2>&1
tail -2 gpurun_out/ncu_gemv.log
