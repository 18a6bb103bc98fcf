set -x
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r3f_tests.txt
timeout 600 python bench.py --steps 10 --no-e2e --no-cpu-baseline --timeline gpurun_out/r3f_timeline.txt > gpurun_out/r3f_bench_n1.json 2> gpurun_out/r3f_bench_n1.err
