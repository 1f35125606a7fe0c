#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_packed_gpu.py tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/bench_multi.py tp > gpurun_out/multi_tp_n2.json 2> gpurun_out/multi_tp_n2.err; tail -1 gpurun_out/multi_tp_n2.json; grep -v "OMP_NUM\|\*\*\*\*" gpurun_out/multi_tp_n2.err | tail -12
