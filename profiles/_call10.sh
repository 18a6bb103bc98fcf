set -x
timeout 900 python -m pytest tests/test_gpu_synth.py tests/test_gpu_parity.py -x -q 2>&1 | tail -8 > gpurun_out/r2j_tests.txt
RFGPU_TEST_SKEW=1 timeout 1500 python -m pytest tests/test_gpu_dist.py -q -rs 2>&1 | tail -12 > gpurun_out/r2j_gpu_dist_n2.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 5 --no-e2e > gpurun_out/r2j_bench_n2.json 2> gpurun_out/r2j_bench_n2.err
tail -5 gpurun_out/r2j_bench_n2.err
timeout 900 python bench.py --steps 5 --no-e2e --no-cpu-baseline > gpurun_out/r2j_bench_n1.json 2> gpurun_out/r2j_bench_n1.err
