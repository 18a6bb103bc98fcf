set -x
timeout 900 python bench.py --steps 5 > gpurun_out/r2h_bench_n1.json 2> gpurun_out/r2h_bench_n1.err
tail -5 gpurun_out/r2h_bench_n1.err
