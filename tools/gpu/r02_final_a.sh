#!/bin/bash
# Round-2 final single-GPU batch: bench lines of every config at HEAD, reference arm, ncu launch list.
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2f_bench_c2.json 2> gpurun_out/r2f_bench_c2.err; tail -c 300 gpurun_out/r2f_bench_c2.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2f_bench_ref_c2.json 2> gpurun_out/r2f_bench_ref_c2.err
timeout 200 python bench.py --config 1 --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/r2f_bench_c1.json 2> gpurun_out/r2f_bench_c1.err
timeout 200 python bench.py --config 3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2f_bench_c3.json 2> gpurun_out/r2f_bench_c3.err
timeout 240 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/r2f_bench_c5.json 2> gpurun_out/r2f_bench_c5.err
for f in c2 ref_c2 c1 c3 c5; do cut -c 1-260 gpurun_out/r2f_bench_$f.json; done
timeout 200 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_f.log 2>&1 && \
  timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/r2f_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_f.log 2>&1
wc -l gpurun_out/r2f_launches.csv
