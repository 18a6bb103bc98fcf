set -x
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5 > gpurun_out/r2f_noalloc_alone.txt
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2f_tests.txt
