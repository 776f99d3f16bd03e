#!/bin/bash
# Round-2 GPU batch I: block-search neighbour tables: parity, full suite, bench, shares.
set -x
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_ints.py -x -q -k "subm_map" 2>&1 | tail -4
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 180 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_c2_i.json 2> gpurun_out/r2_bench_c2_i.err; tail -c 300 gpurun_out/r2_bench_c2_i.err; cut -c 1-330 gpurun_out/r2_bench_c2_i.json
timeout 120 python tools/kernel_shares.py --top 60 2>&1 | grep -E "lookup|stem|blocks|insert|device total"
