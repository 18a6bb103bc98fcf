set -x
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r3j_smoke.txt 2>&1
timeout 900 python bench.py > gpurun_out/r3j_bench_default.json 2> gpurun_out/r3j_bench_default.err
timeout 900 python bench.py --impl reference > gpurun_out/r3j_bench_reference.json 2> gpurun_out/r3j_bench_reference.err
timeout 900 python bench.py --workload optional --triples 500000000 > gpurun_out/r3j_cfg4.json 2> gpurun_out/r3j_cfg4.err
timeout 600 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r3j_b.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r3j_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r3j_ncu.log 2>&1
