set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2c_tests.txt
timeout 300 python bench.py --hash-join --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2c_bench_lookup_unsorted.json 2> gpurun_out/r2c_1.err
timeout 300 python bench.py --hash-join --sorted-build --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2c_bench_lookup_sorted.json 2> gpurun_out/r2c_2.err &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_lookup_join -s 3 -c 2 -o gpurun_out/r2c_lookup python bench.py --hash-join --sorted-build --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r2c_ncu.log 2>&1
