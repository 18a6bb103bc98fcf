set -x
RFGPU_TEST_SKEW=1 timeout 900 python -m pytest tests/test_gpu_dist.py -q -rs -k "8" 2>&1 | tail -8 > gpurun_out/r3k_gpu_dist_n8.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 8 --steps 10 --warmup 3 --timeline gpurun_out/r3k_timeline_n8.txt > gpurun_out/r3k_bench_n8.json 2> gpurun_out/r3k_bench_n8.err
tail -3 gpurun_out/r3k_bench_n8.err
