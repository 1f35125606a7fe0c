#!/bin/bash
timeout 600 python tools/bench_gemv.py 1 2>&1 | grep -v "ring\|1cta\|split(7"
timeout 600 python tools/bench_gemv.py 1 static 2>&1 | grep "inputs\|fused" | grep -v "ring\|1cta"
echo "=== layer (chain)"; timeout 300 python tools/gemv_timeline.py layer 1 2>&1 | tail -8
