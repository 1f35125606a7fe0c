#!/bin/bash
# launch list of the bench command, this library's fake-quant kernels only (per-launch durations under ncu are cold-cache and
# serialised: shares, not absolutes)
set -o pipefail
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"int_thr|int_ste|int_fq" -c 400 --csv --log-file gpurun_out/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench68_ncu.log 2>&1; tail -n 2 gpurun_out/bench68_ncu.log | cut -c1-300
