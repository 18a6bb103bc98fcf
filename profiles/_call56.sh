set -x
timeout 300 python __graft_entry__.py smoke > gpurun_out/r4i_smoke.txt 2>&1; tail -1 gpurun_out/r4i_smoke.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r4i_gpu_tests.txt
tail -3 gpurun_out/r4i_gpu_tests.txt
timeout 300 python profiles/tv_microbench.py > gpurun_out/r4i_tv_microbench.json 2> gpurun_out/r4i_tv_microbench.err
timeout 600 python bench.py > gpurun_out/r4i_bench_default.json 2> gpurun_out/r4i_bench_default.err
tail -2 gpurun_out/r4i_bench_default.err
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r4i_bench_reference.json 2> gpurun_out/r4i_bench_reference.err
timeout 600 python bench.py --workload optional --triples 500000000 > gpurun_out/r4i_cfg4.json 2> gpurun_out/r4i_cfg4.err
tail -2 gpurun_out/r4i_cfg4.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r4i_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r4i_ncu.log 2>&1
tail -2 gpurun_out/r4i_ncu.log
python - <<'P'
import json
d=json.load(open("gpurun_out/r4i_tv_microbench.json"))
for r in d["results"]:
    print(r["operator"][:60], r["kernel_us"], r["frac_of_measured_hbm_peak"], {k:v["us"] for k,v in r["kernels"].items()})
b=json.loads(open("gpurun_out/r4i_bench_default.json").read().strip().splitlines()[-1])
print(b["value"], b["ms_per_step"], b["e2e"]["value"], b["roofline"]["frac"], b["cpu_baseline"]["value"], {k:(round(v["ms_per_launch"]*1e3,1), round(v["frac"] or 0,3)) for k,v in b["config"]["kernels"].items()})
r=json.loads(open("gpurun_out/r4i_bench_reference.json").read().strip().splitlines()[-1])
print("reference", r.get("value"), r.get("ms_per_step"))
P
