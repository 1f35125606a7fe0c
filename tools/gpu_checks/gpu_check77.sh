#!/bin/bash
# 8 GPUs: autograd / transport check of the gather path at world 8, bench.py --gpus 8 with this build
set -o pipefail
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29661 tools/check_shard_gather_autograd.py 2>&1 | grep -E '^\{|^[0-7] \[' | tee gpurun_out/shard_autograd_n8.json | cut -c1-400
timeout 600 $TR --master-port 29662 bench.py --gpus 8 --steps 10 --warmup 3 2>&1 | grep '^{' | tee gpurun_out/bench_n8.json | cut -c1-400
