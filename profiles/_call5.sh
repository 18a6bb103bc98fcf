set -x
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5 > gpurun_out/r2e_noalloc_alone.txt
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2e_tests.txt
nvcc -gencode arch=compute_100a,code=sm_100a -O3 --extended-lambda -o /tmp/hwbench profiles/hwbench.cu && /tmp/hwbench > gpurun_out/r2e_hwbench.json 2> gpurun_out/r2e_hwbench.err
MB_ONLY=sort timeout 600 python profiles/microbench.py > gpurun_out/r2e_sort_microbench.json 2> gpurun_out/r2e_sort_microbench.err
timeout 300 python bench.py --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2e_bench_merge.json 2> gpurun_out/r2e_0.err
timeout 300 python bench.py --hash-join --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2e_bench_lookup_unsorted.json 2> gpurun_out/r2e_1.err
timeout 300 python bench.py --hash-join --sorted-build --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2e_bench_lookup_sorted.json 2> gpurun_out/r2e_2.err
timeout 300 python bench.py --hash-join --no-lookup-join --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2e_bench_hash.json 2> gpurun_out/r2e_3.err
