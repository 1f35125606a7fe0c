#!/bin/bash
# 8 GPUs: quantise-before-all-gather three-way comparison, layer-sharded run_clip check
set -o pipefail
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
echo "=== shard gather n8"; timeout 600 $TR --master-port 29621 tools/bench_shard_gather.py 2>&1 | tee gpurun_out/shard_gather_n8.log | grep '^{' | tee gpurun_out/shard_gather_n8.json
echo "=== run_clip sharded n8"; timeout 600 $TR --master-port 29622 tools/check_run_clip_sharded.py 2>&1 | tee gpurun_out/run_clip_sharded_n8.log | grep '^{' | tee gpurun_out/run_clip_sharded_n8.json
tail -n 3 gpurun_out/shard_gather_n8.log; tail -n 3 gpurun_out/run_clip_sharded_n8.log
