set -x
nvcc -gencode arch=compute_100a,code=sm_100a -O3 --extended-lambda -o /tmp/p2pbench profiles/p2pbench.cu && timeout 120 /tmp/p2pbench > gpurun_out/r2n_p2pbench.json 2> gpurun_out/r2n_p2pbench.err
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "three_and_four or group_count" 2>&1 | tail -3 > gpurun_out/r2n_tests.txt
