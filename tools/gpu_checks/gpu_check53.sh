#!/bin/bash
timeout 300 python tools/prof_gemm.py && timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_w2_kernel --launch-skip 2 -c 1 -o gpurun_out/gemm_mt2_full -f python tools/prof_gemm.py > gpurun_out/gemm_ncu.log 2>&1
tail -2 gpurun_out/gemm_ncu.log
