set -x
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu-baseline --host-profile gpurun_out/r2t_host_profile_full.txt > gpurun_out/r2t_bench_n1.json 2> gpurun_out/r2t_bench_n1.err
timeout 600 python bench.py --triples 2000000 --steps 5 --no-e2e --no-cpu-baseline --host-profile gpurun_out/r2t_host_profile_tiny.txt > gpurun_out/r2t_bench_tiny.json 2> gpurun_out/r2t_bench_tiny.err
