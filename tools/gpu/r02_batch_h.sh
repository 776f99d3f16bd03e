#!/bin/bash
# Round-2 GPU batch H: full GPU suite with tight timeout, then headline bench.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 180 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_c2_h.json 2> gpurun_out/r2_bench_c2_h.err; tail -c 300 gpurun_out/r2_bench_c2_h.err; cut -c 1-330 gpurun_out/r2_bench_c2_h.json
