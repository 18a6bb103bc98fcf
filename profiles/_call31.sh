set -x
export RFGPU_LIB=$PWD/.wip/librfgpu_diag.so
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 3 --no-e2e --timeline gpurun_out/r3e_timeline_diag.txt > gpurun_out/r3e_bench_diag.json 2> gpurun_out/r3e_bench_diag.err
tail -5 gpurun_out/r3e_bench_diag.err
