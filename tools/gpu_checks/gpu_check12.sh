#!/bin/bash
mkdir -p gpurun_out
python tools/prof_gemv.py > gpurun_out/plain_gemv.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemv_w2s -s 4 -c 3 -f -o gpurun_out/prof_gemv_s python tools/prof_gemv.py > gpurun_out/ncu_gemv.log 2>&1
tail -3 gpurun_out/ncu_gemv.log; ls -la gpurun_out/*.ncu-rep
