set -x
timeout 600 python bench.py --steps 5 --no-e2e --no-cpu-baseline --timeline gpurun_out/r2v_timeline.txt > gpurun_out/r2v_bench_n1.json 2> gpurun_out/r2v_bench_n1.err
