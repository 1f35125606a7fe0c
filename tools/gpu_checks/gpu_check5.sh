#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python tools/prof_fq.py > gpurun_out/plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'int_' -c 2 -f -o gpurun_out/prof_thr python tools/prof_fq.py > gpurun_out/ncu_thr.log 2>&1
tail -3 gpurun_out/ncu_thr.log
