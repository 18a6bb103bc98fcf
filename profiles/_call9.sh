set -x
RFGPU_TEST_SKEW=1 timeout 1500 python -m pytest tests/test_gpu_dist.py -q -rs 2>&1 | tail -25 > gpurun_out/r2i_gpu_dist_n2.txt
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 5 --no-e2e > gpurun_out/r2i_bench_n2.json 2> gpurun_out/r2i_bench_n2.err
tail -5 gpurun_out/r2i_bench_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 2 --steps 5 --no-e2e --layout range > gpurun_out/r2i_bench_n2_range.json 2> gpurun_out/r2i_bench_n2_range.err
tail -5 gpurun_out/r2i_bench_n2_range.err
