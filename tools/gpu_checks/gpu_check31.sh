#!/bin/bash
timeout 120 python tools/prof_gemv_big.py 0 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_w2s --launch-skip 1 -c 1 -o gpurun_out/gemv_big_full -f python tools/prof_gemv_big.py 0 > gpurun_out/gemv_big_ncu.log 2>&1
tail -3 gpurun_out/gemv_big_ncu.log
