set -x
timeout 600 python -m pytest tests/test_gpu_zz_typed_value.py tests/test_gpu_zz_dict.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r4a_gpu_tv_tests.txt
tail -5 gpurun_out/r4a_gpu_tv_tests.txt
timeout 400 python profiles/tv_microbench.py > gpurun_out/r4a_tv_microbench.json 2> gpurun_out/r4a_tv_microbench.err
tail -3 gpurun_out/r4a_tv_microbench.err
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r4a_gpu_tests.txt
tail -4 gpurun_out/r4a_gpu_tests.txt
