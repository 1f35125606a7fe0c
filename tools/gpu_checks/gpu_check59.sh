#!/bin/bash
# packed NF3 GEMV second version (pair lookup): parity tests, timing vs the first version, ncu
set -o pipefail
timeout 900 python -m pytest tests/test_nf3_packed_gpu.py -x -q 2>&1 | tail -5
echo "=== v2"; timeout 300 python tools/bench_nf3_gemv.py 2>&1 | tee gpurun_out/nf3_gemv_v2.txt | tail -8
echo "=== v1"; BDQ_NF3_GEMV_V1=1 timeout 300 python tools/bench_nf3_gemv.py 2>&1 | tee gpurun_out/nf3_gemv_v1.txt | tail -8
echo "=== ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:nf3_gemv2 -c 1 -s 30 -o gpurun_out/nf3_gemv2 -f python tools/bench_nf3_gemv.py > gpurun_out/nf3_gemv2_ncu.log 2>&1; tail -2 gpurun_out/nf3_gemv2_ncu.log
