set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2q_tests.txt
timeout 900 python bench.py --steps 10 --no-e2e --no-cpu-baseline > gpurun_out/r2q_bench_n1.json 2> gpurun_out/r2q_bench_n1.err
