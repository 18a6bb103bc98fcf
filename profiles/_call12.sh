set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2l_tests.txt
timeout 900 python bench.py --workload optional > gpurun_out/r2l_cfg4.json 2> gpurun_out/r2l_cfg4.err
tail -5 gpurun_out/r2l_cfg4.err
