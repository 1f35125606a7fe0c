#!/bin/bash
# after factoring reduce_scatter_rows: 1-GPU shard-gather tests + the 2-GPU autograd / transport check
set -o pipefail
timeout 600 python -m pytest tests/test_shard_gather_gpu.py -x -q 2>&1 | tail -n 2
bash tools/gpu_checks/gpu_check66.sh 2>&1 | grep '^{' | head -1
