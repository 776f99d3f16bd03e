#!/bin/bash
# bench lines after the roofline-leg refactor (config 2 headline, config 5 with its roofline)
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench_c2.json 2> gpurun_out/r2h_bench_c2.err; tail -c 300 gpurun_out/r2h_bench_c2.err; cut -c 1-200 gpurun_out/r2h_bench_c2.json
timeout 240 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/r2h_bench_c5.json 2> gpurun_out/r2h_bench_c5.err; tail -c 300 gpurun_out/r2h_bench_c5.err; cut -c 1-200 gpurun_out/r2h_bench_c5.json
