#!/bin/bash
# Round-2 GPU batch F: block-occupancy stem search: parity, then the headline bench and kernel shares.
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_floats.py -x -q -k "fused_stem or forward_golden or subm" 2>&1 | tail -5
python -m pytest tests/test_gpu_parity_fullsize.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_c2_f.json 2> gpurun_out/r2_bench_c2_f.err; tail -c 500 gpurun_out/r2_bench_c2_f.err; cut -c 1-330 gpurun_out/r2_bench_c2_f.json
python tools/kernel_shares.py --top 12 2>&1 | grep -v Warn | tail -16
