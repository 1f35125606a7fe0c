#!/bin/bash
# end-to-end run_clip on a random-init 7B architecture; run_clip parity tests
set -o pipefail
timeout 600 python -m pytest tests/test_run_clip_gpu.py -x -q 2>&1 | tail -n 2
timeout 600 python tools/bench_run_clip_7b.py 128 1024 32 2>&1 | tee gpurun_out/run_clip_7b.log | tail -n 3 | cut -c1-700
grep '^{' gpurun_out/run_clip_7b.log > gpurun_out/run_clip_7b.json
