#!/bin/bash
# N4: quantise-before-all-gather -- parity tests, 1-GPU kernel rooflines, ncu of encode/decode
set -o pipefail
timeout 900 python -m pytest tests/test_shard_gather_gpu.py -x -q 2>&1 | tail -8
echo "=== bench"; timeout 300 python tools/bench_shard_gather.py 2>&1 | tee gpurun_out/shard_gather_n1.json | tail -3
echo "=== ncu"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:int2_ -c 2 -s 4 -o gpurun_out/shard_gather -f python tools/bench_shard_gather.py > gpurun_out/shard_gather_ncu.log 2>&1; tail -2 gpurun_out/shard_gather_ncu.log
