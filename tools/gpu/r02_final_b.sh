#!/bin/bash
# Round-2 closing single-GPU batch at HEAD: bench lines of every config + reference arm, full GPU suite, smoke, config-5
# shares, ncu of the wide conv launches and the ncu launch list of the bench command.  Every command under its own timeout.
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r2g_bench_c2.json 2> gpurun_out/r2g_bench_c2.err; tail -c 300 gpurun_out/r2g_bench_c2.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2g_bench_ref_c2.json 2> gpurun_out/r2g_bench_ref_c2.err
timeout 200 python bench.py --config 1 --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/r2g_bench_c1.json 2> gpurun_out/r2g_bench_c1.err
timeout 200 python bench.py --config 3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2g_bench_c3.json 2> gpurun_out/r2g_bench_c3.err
timeout 240 python bench.py --config 5 --steps 10 --warmup 3 > gpurun_out/r2g_bench_c5.json 2> gpurun_out/r2g_bench_c5.err
for f in c2 ref_c2 c1 c3 c5; do cut -c 1-260 gpurun_out/r2g_bench_$f.json; done
timeout 700 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2g_pytest_tail.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/r2g_smoke_tail.txt
timeout 200 python tools/lin_shares.py > gpurun_out/r2g_lin_shares.txt 2>&1; grep -E "step|rows per|conv launches" gpurun_out/r2g_lin_shares.txt | cut -c 1-600
timeout 120 python tools/conv_bench.py 5 3 > gpurun_out/r2g_conv_plain.txt 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:subm_conv_tc_kernel --launch-skip 27 --launch-count 7 \
    -o gpurun_out/r2g_conv_wide python tools/conv_bench.py 5 3 > gpurun_out/r2g_conv_ncu.log 2>&1
ls -la gpurun_out/r2g_conv_wide.ncu-rep
timeout 200 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_g.log 2>&1 && \
  timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/r2g_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_g.log 2>&1
wc -l gpurun_out/r2g_launches.csv
