#!/bin/bash
timeout 900 python -m pytest tests/test_fakequant_gpu.py -m gpu -x -q 2>&1 | tail -3
echo "=== nf3 prof"; timeout 300 python tools/microbench.py 2>&1 | grep "nf3"
echo "=== bench"; timeout 900 python bench.py > gpurun_out/bench_sec.json 2> gpurun_out/bench_sec.err; tail -1 gpurun_out/bench_sec.json; tail -3 gpurun_out/bench_sec.err
