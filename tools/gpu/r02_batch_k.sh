#!/bin/bash
# LayerNorm C = 768 vector path + bf16 training head: tests, config-5 bench, step shares of configs 2 and 5.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_floats.py -x -q -k "ln_residual" 2>&1 | tail -3
timeout 400 python -m pytest tests/test_gpu_sonata.py -x -q 2>&1 | tail -3
timeout 240 python bench.py --config 5 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/k_bench_c5.json 2> gpurun_out/k_bench_c5.err; cut -c 1-300 gpurun_out/k_bench_c5.json
timeout 200 python tools/lin_shares.py > gpurun_out/k_lin_shares.txt 2>&1; head -30 gpurun_out/k_lin_shares.txt
timeout 200 python tools/kernel_shares.py --top 70 > gpurun_out/k_shares_c2.txt 2>&1; head -75 gpurun_out/k_shares_c2.txt
