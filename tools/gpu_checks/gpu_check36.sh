#!/bin/bash
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/bench_multi.py tp > gpurun_out/multi_tp_n8.json 2> gpurun_out/multi_tp_n8.err; tail -1 gpurun_out/multi_tp_n8.json; grep -v "OMP_NUM\|\*\*\*\*" gpurun_out/multi_tp_n8.err | tail -8
