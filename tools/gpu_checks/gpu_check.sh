#!/bin/bash
# One GPU-box pass: parity tests, micro-benchmark, headline bench, ncu launch list + full captures.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python tools/microbench.py > gpurun_out/microbench.log 2>&1; tail -20 gpurun_out/microbench.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -2 gpurun_out/bench_ref.json
python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python tools/prof_fq.py > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'int_|nf3_' -c 8 -f -o gpurun_out/prof_fq python tools/prof_fq.py > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
