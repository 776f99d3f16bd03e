#!/bin/bash
# Round-2 GPU batch G: proj + LayerNorm + MLP fused kernel: model parity, headline bench, shares.  Tight timeouts.
set -x
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_floats.py -x -q -k "native_block or forward_golden or mlp_fused" 2>&1 | tail -4
timeout 180 python -m pytest tests/test_gpu_parity_fullsize.py tests/test_gpu_sonata.py -x -q 2>&1 | tail -3
timeout 180 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_bench_c2_g.json 2> gpurun_out/r2_bench_c2_g.err; tail -c 500 gpurun_out/r2_bench_c2_g.err; cut -c 1-330 gpurun_out/r2_bench_c2_g.json
timeout 120 python tools/kernel_shares.py --top 14 2>&1 | grep -v -i warn | head -18
