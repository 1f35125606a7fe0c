#!/bin/bash
timeout 600 python tools/gemv_variants.py 1 3 35 51 2>&1 | tail -6
echo "=== big"; timeout 300 python tools/gemv_big.py 1 2>&1 | tail -5
timeout 300 python tools/gemv_big.py 17 2>&1 | tail -5
echo "=== timeline 35"; timeout 300 python tools/gemv_timeline.py layer 35 pf 2>&1 | tail -9
