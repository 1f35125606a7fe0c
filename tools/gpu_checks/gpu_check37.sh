#!/bin/bash
set -o pipefail
date +%s > gpurun_out/t0
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "=== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
echo "=== bench"; /usr/bin/time -v timeout 1200 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; tail -c 1500 gpurun_out/bench_final.json; grep -i "elapsed\|error" gpurun_out/bench_final.err | tail -3
echo "=== ref arm"; timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 600 gpurun_out/bench_ref.json
