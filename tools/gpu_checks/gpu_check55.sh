#!/bin/bash
# Re-entry baseline: the whole GPU suite, smoke, both bench arms
set -o pipefail
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
echo "=== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
echo "=== bench"; timeout 600 python bench.py 2>&1 | tee gpurun_out/bench55.json | tail -2
echo "=== ref"; timeout 600 python bench.py --impl reference 2>&1 | tee gpurun_out/bench55_ref.json | tail -2
