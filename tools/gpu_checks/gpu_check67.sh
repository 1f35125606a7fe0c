#!/bin/bash
# GPTQ export with the threshold predicates spread over the group lanes (slower: 44 vs 41 us, reverted); group autograd test
set -o pipefail
timeout 900 python -m pytest tests/test_gptq_gpu.py tests/test_shard_gather_gpu.py -x -q 2>&1 | tail -4
echo "=== bench"; timeout 300 python tools/bench_gptq.py 2>&1 | tee gpurun_out/gptq_bench.txt | tail -8
echo "=== ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:gptq -c 1 -s 12 -o gpurun_out/gptq_export -f python tools/bench_gptq.py > gpurun_out/gptq_ncu.log 2>&1; tail -n 2 gpurun_out/gptq_ncu.log
