#!/bin/bash
# 8-GPU pass: headline bench at N=8, sharded autoclip, TP (NCCL + fused epilogues)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; tail -1 gpurun_out/bench_n8.json | cut -c1-300; grep -v "OMP_NUM\|\*\*\*\*" gpurun_out/bench_n8.err | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 tools/bench_multi.py all > gpurun_out/multi_n8.json 2> gpurun_out/multi_n8.err; tail -1 gpurun_out/multi_n8.json; grep -v "OMP_NUM\|\*\*\*\*" gpurun_out/multi_n8.err | tail -8
