set -x
export RFGPU_LIB=$PWD/.wip/librfgpu_s4.so
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 10 --no-e2e > gpurun_out/r3l_bench_n2_s4.json 2> gpurun_out/r3l_bench_n2_s4.err
tail -3 gpurun_out/r3l_bench_n2_s4.err
