#!/bin/bash
# (historical) two-row-tile variant of the tcgen05 conv: parity, micro-bench per mode, config 2 / 5 effect.  The variant was
# measured and reverted (profiles/r02_fusions.md); CONV_ASYNC bits 1-2 no longer select anything at HEAD.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_floats.py -x -q -k "subm_conv" 2>&1 | tail -4
CONV_ASYNC=5 timeout 120 python tools/conv_bench.py 5 10 > gpurun_out/m_conv_r2.txt 2>&1; grep stage gpurun_out/m_conv_r2.txt | tail -4
PTV3_CONV_ASYNC=3 timeout 200 python tools/lin_shares.py > gpurun_out/m_lin_shares_r1.txt 2>&1; grep -E "step|subm_conv|conv launches" gpurun_out/m_lin_shares_r1.txt | cut -c 1-1500
PTV3_CONV_ASYNC=5 timeout 200 python tools/lin_shares.py > gpurun_out/m_lin_shares_r2.txt 2>&1; grep -E "step|subm_conv|conv launches" gpurun_out/m_lin_shares_r2.txt | cut -c 1-1500
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/m_bench_c2.json 2> gpurun_out/m_bench_c2.err; cut -c 1-200 gpurun_out/m_bench_c2.json
