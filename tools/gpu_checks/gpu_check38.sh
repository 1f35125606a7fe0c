#!/bin/bash
s=$(date +%s)
timeout 1200 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "rc=$? wall=$(( $(date +%s) - s )) s"; tail -c 2500 gpurun_out/bench_final.json; tail -3 gpurun_out/bench_final.err
