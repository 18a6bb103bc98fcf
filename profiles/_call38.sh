set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r3n_bench_n4.json 2> gpurun_out/r3n_bench_n4.err
tail -3 gpurun_out/r3n_bench_n4.err
