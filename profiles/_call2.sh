set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_minus.py tests/test_gpu_zy_plan_shapes.py -x -q 2>&1 | tail -15 > gpurun_out/r2b_tests.txt
timeout 300 python bench.py --hash-join --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2b_bench_lookup_unsorted.json 2> gpurun_out/r2b_1.err
timeout 300 python bench.py --hash-join --sorted-build --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2b_bench_lookup_sorted.json 2> gpurun_out/r2b_2.err
