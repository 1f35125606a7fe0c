#!/bin/bash
mkdir -p gpurun_out
for s in gate q down layer; do
  echo "=== $s flags=1 (pdl)"; timeout 300 python tools/gemv_timeline.py $s 1 2>&1 | tail -30
done
echo "=== gate flags=0 (no pdl)"; timeout 300 python tools/gemv_timeline.py gate 0 2>&1 | tail -12
echo "=== gate flags=3 (pdl+2cta)"; timeout 300 python tools/gemv_timeline.py gate 3 2>&1 | tail -12
