set -x
RFGPU_TEST_SKEW=1 timeout 2400 python -m pytest tests/test_gpu_dist.py -q -rs -k "8 or 4 or overflow" 2>&1 | tail -12 > gpurun_out/r2k_gpu_dist_n4_n8.txt
for N in 8 4; do
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N bench.py --gpus $N --steps 10 > gpurun_out/r2k_bench_n$N.json 2> gpurun_out/r2k_bench_n$N.err
tail -3 gpurun_out/r2k_bench_n$N.err
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 8 --steps 5 --no-e2e --layout range > gpurun_out/r2k_bench_n8_range.json 2> gpurun_out/r2k_bench_n8_range.err
nvidia-smi topo -m > gpurun_out/r2k_topo.txt 2>&1; nproc >> gpurun_out/r2k_topo.txt
