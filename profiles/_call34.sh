set -x
export RFGPU_LIB=$PWD/.wip/librfgpu_diag2.so
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 3 --no-e2e --timeline gpurun_out/r3i_timeline_diag2.txt > gpurun_out/r3i_bench_diag2.json 2> gpurun_out/r3i_bench_diag2.err
tail -3 gpurun_out/r3i_bench_diag2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 2 --steps 3 --no-e2e --no-bulk-shuffle --timeline gpurun_out/r3i_timeline_diag2_nobulk.txt > gpurun_out/r3i_bench_diag2_nobulk.json 2> gpurun_out/r3i_bench_diag2_nobulk.err
