set -x
nvcc -gencode arch=compute_100a,code=sm_100a -O3 --extended-lambda -o /tmp/hwbench profiles/hwbench.cu && /tmp/hwbench > gpurun_out/r2a_hwbench.json 2> gpurun_out/r2a_hwbench.err
MB_ONLY=sort timeout 600 python profiles/microbench.py > gpurun_out/r2a_sort_microbench.json 2> gpurun_out/r2a_sort_microbench.err
timeout 300 python bench.py --hash-join --sort-merge-join --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2a_bench_sm.json 2> gpurun_out/r2a_bench_sm.err
timeout 300 python bench.py --hash-join --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2a_bench_hash.json 2> gpurun_out/r2a_bench_hash.err
timeout 300 python bench.py --device-build --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/r2a_bench_devbuild.json 2> gpurun_out/r2a_bench_devbuild.err
nvidia-smi topo -m > gpurun_out/r2a_topo.txt 2>&1
nproc > gpurun_out/r2a_nproc.txt; free -g >> gpurun_out/r2a_nproc.txt; lscpu | head -30 >> gpurun_out/r2a_nproc.txt
