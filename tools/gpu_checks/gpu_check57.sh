#!/bin/bash
# N4 on 2 GPUs: shard-gather + run_clip tests, encode kernel after the IMAD-Horner change, 2-GPU three-way comparison
set -o pipefail
timeout 900 python -m pytest tests/test_shard_gather_gpu.py tests/test_run_clip_gpu.py -x -q 2>&1 | tail -8
echo "=== n1"; timeout 300 python tools/bench_shard_gather.py 2>&1 | tee gpurun_out/shard_gather_n1.json | tail -2
echo "=== n2"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tools/bench_shard_gather.py 2>&1 | tee gpurun_out/shard_gather_n2.json | tail -5
