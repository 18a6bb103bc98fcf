set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2o_tests.txt
timeout 900 python bench.py --steps 10 --no-e2e --no-cpu-baseline > gpurun_out/r2o_bench_n1.json 2> gpurun_out/r2o_bench_n1.err
MB_ONLY=semi MB_TRIPLES=20000000 timeout 300 python profiles/microbench.py > gpurun_out/r2o_semi_microbench.json 2> gpurun_out/r2o_semi.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2o_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_filter_compact|k_lookup_join|k_merge_join|k_dense_heads_compact" -s 12 -c 8 -o gpurun_out/r2o_kernels python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2o_ncu.log 2>&1
