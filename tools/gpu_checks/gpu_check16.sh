#!/bin/bash
# 2-GPU pass: sharded autoclip + TP GEMV/GEMM, then the headline bench at N=2
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/bench_multi.py all > gpurun_out/multi_n2.json 2> gpurun_out/multi_n2.err; tail -2 gpurun_out/multi_n2.json; tail -5 gpurun_out/multi_n2.err
timeout 600 python tools/bench_multi.py autoclip > gpurun_out/multi_n1.json 2> gpurun_out/multi_n1.err; tail -1 gpurun_out/multi_n1.json; tail -3 gpurun_out/multi_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -1 gpurun_out/bench_n2.json | cut -c1-400; tail -3 gpurun_out/bench_n2.err
