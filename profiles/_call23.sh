set -x
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2w_tests.txt
timeout 600 python bench.py --steps 10 --no-e2e --no-cpu-baseline --timeline gpurun_out/r2w_timeline.txt > gpurun_out/r2w_bench_n1.json 2> gpurun_out/r2w_bench_n1.err
