#!/bin/bash
# Round-2 final scaling batch (one 8-GPU box): weak headline at N = 8, strong config 3 at N = 8 and N = 4.
set -x
mkdir -p gpurun_out
run() {  # n config scaling tag
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29600 + $1)) \
    bench.py --gpus $1 --config $2 --scaling $3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2f_scale_$4.json 2> gpurun_out/r2f_scale_$4.err
  tail -c 300 gpurun_out/r2f_scale_$4.err
}
run 8 2 weak c2_weak_n8
run 8 3 strong c3_strong_n8
run 4 3 strong c3_strong_n4
run 2 3 strong c3_strong_n2
cat gpurun_out/r2f_scale_*.json | cut -c 1-200
