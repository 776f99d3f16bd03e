#!/bin/bash
# Round-2 GPU batch D: the whole GPU suite after the PT-v3m2 / DropPath changes, then configs[4] and the headline again.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 600 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/r2_bench_c5.json 2> gpurun_out/r2_bench_c5.err; tail -c 800 gpurun_out/r2_bench_c5.err; cut -c 1-900 gpurun_out/r2_bench_c5.json
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_c2_d.json 2> gpurun_out/r2_bench_c2_d.err; cut -c 1-300 gpurun_out/r2_bench_c2_d.json
python tools/lin_shares.py 2>&1 | tail -32 > gpurun_out/r2_lin_shares.txt; head -12 gpurun_out/r2_lin_shares.txt
