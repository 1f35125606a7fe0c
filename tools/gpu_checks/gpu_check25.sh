#!/bin/bash
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nf3_thr_fwd -c 1 -o gpurun_out/nf3_thr_full -f python tools/prof_fq.py > gpurun_out/nf3_ncu.log 2>&1
tail -3 gpurun_out/nf3_ncu.log
