set -x
timeout 900 python -m pytest tests/test_gpu_synth.py -x -q 2>&1 | tail -15 > gpurun_out/r2g_synth_tests.txt
