set -x
python -m pytest tests/test_gpu_parity_fullsize.py "tests/test_gpu_train.py::test_dense_branch_two_scene_batch_golden" tests/test_gpu_floats.py::test_forward_golden -m gpu -q -s 2>&1 | grep -E "config-|batch-16|dense-branch|scene [01] |passed|failed" > gpurun_out/r2_parity_numbers.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_bench_ref_c2.json 2>> gpurun_out/r2_bench_c2.err
python bench.py --config 1 --steps 50 --warmup 10 > gpurun_out/r2_bench_c1.json 2> gpurun_out/r2_bench_c1.err
python bench.py --config 3 --steps 20 --warmup 5 > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2_bench_c3.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
tail -2 gpurun_out/r2_bench_c2.err gpurun_out/r2_bench_c1.err gpurun_out/r2_bench_c3.err
cat gpurun_out/r2_parity_numbers.log
