#!/bin/bash
# 2 GPUs: quantized_all_gather autograd (forward bit-equal, shard gradient = rows of the full-weight STE backward of the summed gradient)
set -o pipefail
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29651 tools/check_shard_gather_autograd.py 2>&1 | tee gpurun_out/shard_autograd_n2.log | grep '^{' | tee gpurun_out/shard_autograd_n2.json
tail -n 5 gpurun_out/shard_autograd_n2.log
