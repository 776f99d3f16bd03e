#!/bin/bash
# Round-2 scaling batch (one 8-GPU box): strong scaling of config 3 (8 scenes in total) at N = 1, 2, 4, 8 and the weak
# headline (config 2, 16 scenes per GPU) at N = 8.  Output lines -> gpurun_out/r2_scale_*.json
set -x
mkdir -p gpurun_out
run() {  # n config scaling tag
  if [ "$1" = 1 ]; then
    timeout 400 python bench.py --gpus 1 --config $2 --scaling $3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_scale_$4.json 2> gpurun_out/r2_scale_$4.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + $1)) \
      bench.py --gpus $1 --config $2 --scaling $3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2_scale_$4.json 2> gpurun_out/r2_scale_$4.err
  fi
  tail -c 600 gpurun_out/r2_scale_$4.err
}
run 1 3 strong c3_strong_n1
run 2 3 strong c3_strong_n2
run 4 3 strong c3_strong_n4
run 8 3 strong c3_strong_n8
run 8 2 weak c2_weak_n8
cat gpurun_out/r2_scale_*.json | cut -c 1-400
