#!/bin/bash
# (historical) two-row-tile conv variant on a 3-stage ring filled by 3 groups: parity + effect.  Reverted (profiles/r02_fusions.md).
set -x
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_floats.py -x -q -k "two_row_tiles or many_tiles" 2>&1 | tail -4
CONV_ASYNC=5 timeout 60 python tools/conv_bench.py 5 10 > gpurun_out/n_conv_r2.txt 2>&1; grep stage gpurun_out/n_conv_r2.txt | tail -3
PTV3_CONV_ASYNC=5 timeout 90 python tools/lin_shares.py > gpurun_out/n_lin_shares_r2.txt 2>&1; grep -E "step|subm_conv|conv launches" gpurun_out/n_lin_shares_r2.txt | cut -c 1-1500
