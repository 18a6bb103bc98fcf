set -x
timeout 70 ncu --set full --clock-control none --import-source on -k regex:k_filter_compact -s 3 -c 2 -o gpurun_out/r4n_filter python bench.py --parts filter --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r4n_ncu.log 2>&1
tail -2 gpurun_out/r4n_ncu.log
ncu -i gpurun_out/r4n_filter.ncu-rep --page raw --csv > gpurun_out/r4n_filter_raw.csv 2>&1
