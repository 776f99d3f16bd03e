#!/bin/bash
# Round-2 GPU batch C: PT-v3m2 tests after the frozen fused path, then BASELINE configs[4] (bench.py --config 5).
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_sonata.py -x -q -s 2>&1 | tee gpurun_out/r2_sonata_tests.log | tail -15
timeout 600 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/r2_bench_c5.json 2> gpurun_out/r2_bench_c5.err; tail -c 1500 gpurun_out/r2_bench_c5.err; cat gpurun_out/r2_bench_c5.json
