#!/bin/bash
timeout 300 python tools/prof_autoclip.py 1024 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:autoclip_kernel --launch-skip 1 -c 1 -o gpurun_out/autoclip_full -f python tools/prof_autoclip.py 1024 > gpurun_out/autoclip_ncu.log 2>&1
tail -2 gpurun_out/autoclip_ncu.log
