set -x
timeout 600 python bench.py --parts star --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2y_star.json 2> gpurun_out/r2y_star.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_merge_join -s 3 -c 1 -o gpurun_out/r2y_merge -f python bench.py --parts star --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r2y_ncu.log 2>&1
