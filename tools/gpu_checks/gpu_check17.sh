#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --layers 4 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --layers 4 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python tools/prof_fq.py > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'int_thr' -c 4 -f -o gpurun_out/prof_fq_flat python tools/prof_fq.py > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log; ls -la gpurun_out/*.csv gpurun_out/*.ncu-rep
