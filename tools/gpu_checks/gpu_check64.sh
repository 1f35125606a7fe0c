#!/bin/bash
# ShardGatherGroup: 1-GPU test, 2-GPU four-way comparison
set -o pipefail
timeout 900 python -m pytest tests/test_shard_gather_gpu.py -x -q 2>&1 | tail -3
echo "=== n2"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 tools/bench_shard_gather.py 2>&1 | tee gpurun_out/shard_gather_n2.log | grep '^{' | tee gpurun_out/shard_gather_n2.json
tail -n 4 gpurun_out/shard_gather_n2.log
