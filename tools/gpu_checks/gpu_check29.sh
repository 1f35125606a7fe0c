#!/bin/bash
echo "=== flags 1"; timeout 300 python tools/gemv_timeline.py layer 1 pf 2>&1 | tail -14
echo "=== flags 3"; timeout 300 python tools/gemv_timeline.py layer 3 pf 2>&1 | tail -14
echo "=== flags 19"; timeout 300 python tools/gemv_timeline.py layer 19 pf 2>&1 | tail -14
