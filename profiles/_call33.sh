set -x
RFGPU_TEST_SKEW=1 timeout 1500 python -m pytest tests/test_gpu_dist.py -q -rs 2>&1 | tail -6 > gpurun_out/r3g_gpu_dist_n2.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 10 --no-e2e --timeline gpurun_out/r3g_timeline_n2.txt > gpurun_out/r3g_bench_n2.json 2> gpurun_out/r3g_bench_n2.err
tail -3 gpurun_out/r3g_bench_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 2 --steps 10 --no-e2e --no-bulk-shuffle > gpurun_out/r3g_bench_n2_nobulk.json 2> gpurun_out/r3g_bench_n2_nobulk.err
