#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_packed_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python tools/bench_gemv.py 1 2>&1 | grep -v "^batch 1 ring" 
for s in gate q; do for pf in nopf pf; do
  echo "=== $s flags=1 (pdl) $pf"; timeout 300 python tools/gemv_timeline.py $s 1 $pf 2>&1 | tail -3; done
done
echo "=== layer"; timeout 300 python tools/gemv_timeline.py layer 1 pf 2>&1 | tail -8
