#!/bin/bash
# full GPU suite + GPTQ export bench/ncu (numbers lost with the previous container) + shard-gather ncu after the IMAD-Horner change
set -o pipefail
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "=== gptq bench"; timeout 300 python tools/bench_gptq.py 2>&1 | tee gpurun_out/gptq_bench.txt | tail -12
echo "=== gptq ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:gptq -c 2 -s 7 -o gpurun_out/gptq_export -f python tools/bench_gptq.py > gpurun_out/gptq_ncu.log 2>&1; tail -2 gpurun_out/gptq_ncu.log
echo "=== shard ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:int2_ -c 2 -s 4 -o gpurun_out/shard_gather -f python tools/bench_shard_gather.py > gpurun_out/shard_gather_ncu.log 2>&1; tail -2 gpurun_out/shard_gather_ncu.log
