#!/bin/bash
# 8 GPUs: quantise-before-all-gather four-way comparison with the one-launch layer group
set -o pipefail
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29641 tools/bench_shard_gather.py 2>&1 | tee gpurun_out/shard_gather_n8.log | grep '^{' | tee gpurun_out/shard_gather_n8.json
tail -n 3 gpurun_out/shard_gather_n8.log
