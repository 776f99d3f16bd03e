#!/bin/bash
# Round-2 GPU batch B: the PT-v3m2 path (grid pooling, forward / training goldens, 768-channel conv).
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_sonata.py -x -q -s 2>&1 | tee gpurun_out/r2_sonata_tests.log | tail -40
python -m pytest tests/test_gpu_floats.py -x -q -k "tcgen05_vs_oracle or native_block or unpool" 2>&1 | tail -5
