set -x
timeout 600 python -m pytest tests/test_gpu_zz_typed_value.py tests/test_gpu_zz_dict.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r4b_gpu_tv_tests.txt
tail -5 gpurun_out/r4b_gpu_tv_tests.txt
timeout 400 python profiles/tv_microbench.py > gpurun_out/r4b_tv_microbench.json 2> gpurun_out/r4b_tv_microbench.err
tail -3 gpurun_out/r4b_tv_microbench.err
TV_ROWS=20000000 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tv -c 8 -o gpurun_out/r4b_tv python profiles/tv_microbench.py > gpurun_out/r4b_ncu.log 2>&1
tail -3 gpurun_out/r4b_ncu.log
ncu -i gpurun_out/r4b_tv.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,lts__t_sector_hit_rate.pct,smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio,smsp__average_warp_latency_issue_stalled_lg_throttle.ratio > gpurun_out/r4b_tv_ncu.csv 2>&1
