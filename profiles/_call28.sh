set -x
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 10 --no-e2e --timeline gpurun_out/r3b_timeline_n2.txt > gpurun_out/r3b_bench_n2.json 2> gpurun_out/r3b_bench_n2.err
tail -5 gpurun_out/r3b_bench_n2.err
