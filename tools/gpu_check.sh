#!/bin/bash
# One entry point for everything that is sent to the B200 box through gpurun (replaces the per-call gpu_checkNN.sh log of
# round 1).  Run from the repository root:   gpurun --timeout 1800 -- 'bash tools/gpu_check.sh <sub-command> [args]'
# Outputs land in gpurun_out/ (merged back by gpurun); copy what should be judged into profiles/.
set -u
mkdir -p gpurun_out
cmd=${1:-help}; shift || true
case "$cmd" in
  tests)        # full GPU parity suite (optionally: a pytest -k expression)
    python -m pytest tests -m gpu -q -p no:cacheprovider ${1:+-k "$1"} 2>&1 | tee gpurun_out/pytest_gpu.log | tail -15 ;;
  bench)        # the driver's N=1 line (+ secondary) and the reference arm
    python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 2500 gpurun_out/bench_n1.json
    python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench_n1.err; cat gpurun_out/bench_ref.json ;;
  bench-n)      # sharded QAT step on N GPUs (gpurun --gpus N): bash tools/gpu_check.sh bench-n 2
    n=${1:-2}
    python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus "$n" \
      --steps 5 --warmup 3 > "gpurun_out/bench_n$n.json" 2> "gpurun_out/bench_n$n.err"; tail -c 3000 "gpurun_out/bench_n$n.json"; tail -3 "gpurun_out/bench_n$n.err" ;;
  launches)     # ncu launch list of the bench command: the 128 bdq:: launches of the 2 timed steps (3 warm-up steps skipped)
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:int_thr" -s 192 -c 128 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; tail -3 gpurun_out/ncu.log ;;
  ncu)          # one --set full capture: bash tools/gpu_check.sh ncu <kernel regex> <out name> <python script + args>
    pat=$1; out=$2; shift 2
    python "$@" > gpurun_out/plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k "regex:$pat" -s 1 -c 1 -o "gpurun_out/$out" python "$@" > gpurun_out/ncu.log 2>&1
    tail -3 gpurun_out/ncu.log ;;
  micro)        # per-kernel micro-benchmarks used in DESIGN.md section 7
    for t in microbench bench_nf3 bench_autoclip bench_gemm bench_gptq gemv_big bench_vs_reference; do
      echo "== tools/$t.py"; timeout 600 python "tools/$t.py" 2>&1 | tail -25; done | tee gpurun_out/micro.log ;;
  gemv)         python tools/bench_gemv.py 1 2>&1 | tee gpurun_out/gemv_b1.log | tail -30 ;;
  gemm)         # prefill GEMM per shape (graph-timed, vs cuBLAS on dequantised weights); BDQ_GEMM_TILE=<MT><NT/128> forces a tile
    python tools/bench_gemm.py 2>&1 | tee gpurun_out/gemm_tiles.txt; python tools/gemm_fixed_cost.py 2>&1 | tee -a gpurun_out/gemm_tiles.txt ;;
  nf3-gemv)     # packed NF3 decode GEMV: streaming kernel, then the pair-lookup kernel it replaced
    python tools/bench_nf3_gemv.py 2>&1 | tee gpurun_out/nf3_gemv_stream.txt; BDQ_NF3_GEMV_STREAM=0 python tools/bench_nf3_gemv.py 2>&1 | tee gpurun_out/nf3_gemv_pair.txt ;;
  g64)          # the train scripts' group size: fake-quant kernels at the bench launch shape + the clip search
    python tools/bench_nf3.py 64 2>&1 | tee gpurun_out/g64.txt; python tools/bench_autoclip.py int 64 2>&1 | tail -1 | tee -a gpurun_out/g64.txt; python tools/bench_autoclip.py nf3 64 2>&1 | tail -1 | tee -a gpurun_out/g64.txt ;;
  ubench)       # instruction-rate microbenchmarks (compiled on the box)
    nvcc -arch=sm_100a -o /tmp/alu_rate tools/ubench/alu_rate.cu && /tmp/alu_rate | tee gpurun_out/alu_rate.txt ;;
  chain)        python tools/chain_timeline.py 2>&1 | tee gpurun_out/chain_timeline.txt ;;
  autoclip-roles) python tools/prof_autoclip_roles.py int 2>&1 | tee gpurun_out/autoclip_roles.txt; python tools/prof_autoclip_roles.py nf3 2>&1 | tee -a gpurun_out/autoclip_roles.txt ;;
  multi)        # multi-GPU checks that are not pytest (gpurun --gpus N): shard-gather autograd, layer-sharded run_clip, TP
    n=${1:-2}
    for t in check_shard_gather_autograd check_run_clip_sharded bench_shard_gather bench_multi bench_push; do
      echo "== tools/$t.py"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 \
        --master-port 29513 "tools/$t.py" 2>&1 | grep -v "^\*\|OMP_NUM" | tail -20; done | tee "gpurun_out/multi_n$n.log" ;;
  *) sed -n 2,6p "$0"; grep -E '^  [a-z-]+\)' "$0" | sed 's/)\s*#/  --/' ;;
esac
