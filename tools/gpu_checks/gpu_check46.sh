#!/bin/bash
timeout 600 python tools/bench_nf3_gemv.py 2>&1 | tail -5
