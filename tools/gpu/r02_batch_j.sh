#!/bin/bash
# Round-2 GPU batch J: streamed-weight fused MLP at C = 128: full suite, headline bench, config 5, shares.
set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 180 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_c2_j.json 2> gpurun_out/r2_bench_c2_j.err; tail -c 300 gpurun_out/r2_bench_c2_j.err; cut -c 1-330 gpurun_out/r2_bench_c2_j.json
timeout 120 python tools/kernel_shares.py --top 60 2>&1 | grep -E "mlp|gelu|device total|ln_residual"
timeout 240 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/r2_bench_c5_j.json 2> gpurun_out/r2_bench_c5_j.err; tail -c 300 gpurun_out/r2_bench_c5_j.err; cut -c 1-250 gpurun_out/r2_bench_c5_j.json
