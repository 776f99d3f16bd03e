#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_floats.py -x -q -k "gelu or affine or mlp or unpool" 2>&1 | tail -3
timeout 400 python -m pytest tests/test_gpu_sonata.py tests/test_gpu_parity_fullsize.py -x -q 2>&1 | tail -3
timeout 240 python bench.py --config 5 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/l_bench_c5.json 2> gpurun_out/l_bench_c5.err; cut -c 1-300 gpurun_out/l_bench_c5.json
timeout 200 python tools/lin_shares.py > gpurun_out/l_lin_shares.txt 2>&1; head -12 gpurun_out/l_lin_shares.txt
