#!/bin/bash
for w in 6 8 12 16 32; do echo "=== waves $w"; BDQ_FQ_WAVES=$w timeout 300 python tools/bench_nf3.py 2>&1 | tail -5; done
