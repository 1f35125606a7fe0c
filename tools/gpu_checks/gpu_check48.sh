#!/bin/bash
timeout 900 python -m pytest tests/test_autoclip_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python tools/bench_multi.py autoclip 2>/dev/null | tail -1
