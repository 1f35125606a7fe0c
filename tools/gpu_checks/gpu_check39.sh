#!/bin/bash
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nf3_thr -c 2 -o gpurun_out/nf3_thr2_full -f python tools/prof_fq.py > gpurun_out/nf3_ncu2.log 2>&1
tail -2 gpurun_out/nf3_ncu2.log
