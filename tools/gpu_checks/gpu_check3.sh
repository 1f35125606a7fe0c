#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python tools/prof_gemv.py > gpurun_out/plain3.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_w2 -s 11 -c 2 -f -o gpurun_out/prof_gemv2 python tools/prof_gemv.py > gpurun_out/ncu_gemv.log 2>&1
tail -5 gpurun_out/ncu_gemv.log
