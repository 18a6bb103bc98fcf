set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_w3c.py -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r2x_tests.txt
timeout 600 python bench.py --steps 10 --no-e2e --no-cpu-baseline > gpurun_out/r2x_bench_n1.json 2> gpurun_out/r2x_bench_n1.err
RFGPU_LIB=$PWD/.wip/librfgpu_nopf.so timeout 600 python bench.py --steps 10 --no-e2e --no-cpu-baseline > gpurun_out/r2x_bench_n1_nopf.json 2> gpurun_out/r2x_bench_n1_nopf.err
