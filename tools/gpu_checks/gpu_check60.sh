#!/bin/bash
# packed NF3 GEMV second version with K split over the warps of a CTA: parity, timing, ncu of the last (lm_head-sized) launch
set -o pipefail
timeout 900 python -m pytest tests/test_nf3_packed_gpu.py -x -q 2>&1 | tail -3
echo "=== v2"; timeout 300 python tools/bench_nf3_gemv.py 2>&1 | tee gpurun_out/nf3_gemv_v2.txt | tail -8
echo "=== ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:nf3_gemv2 --kernel-id :::900 -o gpurun_out/nf3_gemv2 -f python tools/bench_nf3_gemv.py > gpurun_out/nf3_gemv2_ncu.log 2>&1; tail -2 gpurun_out/nf3_gemv2_ncu.log
