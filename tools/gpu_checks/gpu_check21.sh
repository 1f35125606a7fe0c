#!/bin/bash
timeout 900 python -m pytest tests/test_fakequant_gpu.py tests/test_autoclip_gpu.py -m gpu -x -q 2>&1 | tail -4
echo "== thr"; timeout 300 python tools/microbench.py 2>&1 | grep "nf3"
echo "== old"; BDQ_NF3_THR=0 timeout 300 python tools/microbench.py 2>&1 | grep "nf3"
