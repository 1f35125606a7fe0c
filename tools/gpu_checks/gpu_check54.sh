#!/bin/bash
# N3: GPTQ-v2 export kernel -- parity tests, timing, launch list
set -o pipefail
timeout 900 python -m pytest tests/test_gptq_gpu.py -x -q 2>&1 | tail -6
echo "=== bench"; timeout 300 python tools/bench_gptq.py 2>&1 | tee gpurun_out/gptq_bench.txt | tail -8
echo "=== ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:gptq_pack -c 2 -s 7 -o gpurun_out/gptq_export -f python tools/bench_gptq.py > gpurun_out/gptq_ncu.log 2>&1; tail -2 gpurun_out/gptq_ncu.log
