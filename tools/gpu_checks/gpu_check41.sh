#!/bin/bash
timeout 900 python -m pytest tests/test_fakequant_gpu.py tests/test_autoclip_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/bench_nf3.py 2>&1 | tail -5
timeout 300 python tools/microbench.py 2>&1 | tail -14
timeout 900 python bench.py --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['frac_of_hbm_peak_per_gpu'], d['clocks'])"
