#!/bin/bash
# BASELINE configs[4] at its own scale at the end of round 2: batch 64 = 8 sweeps per GPU x 8 GPUs.
set -x
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus 8 --config 5 --steps 10 --warmup 3 > gpurun_out/r2g_scale_c5_n8.json 2> gpurun_out/r2g_scale_c5_n8.err
tail -c 400 gpurun_out/r2g_scale_c5_n8.err; cut -c 1-400 gpurun_out/r2g_scale_c5_n8.json
