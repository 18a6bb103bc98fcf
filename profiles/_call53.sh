set -x
timeout 600 python -m pytest tests/test_gpu_zz_typed_value.py tests/test_gpu_zz_dict.py -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r4f_gpu_tv_tests.txt
cat gpurun_out/r4f_gpu_tv_tests.txt
timeout 400 python profiles/tv_microbench.py > gpurun_out/r4f_tv_microbench.json 2> gpurun_out/r4f_tv_microbench.err
tail -3 gpurun_out/r4f_tv_microbench.err
timeout 600 python bench.py > gpurun_out/r4f_bench_default.json 2> gpurun_out/r4f_bench_default.err
tail -3 gpurun_out/r4f_bench_default.err
python - <<'P'
import json
d=json.load(open("gpurun_out/r4f_tv_microbench.json"))
for r in d["results"]:
    print(r["operator"][:60], r["kernel_us"], r["frac_of_measured_hbm_peak"], {k:v["us"] for k,v in r["kernels"].items()})
b=json.loads(open("gpurun_out/r4f_bench_default.json").read().strip().splitlines()[-1])
print(b["value"], b["ms_per_step"], b["e2e"], {k:(round(v["ms_per_launch"]*1e3,1), round(v["frac"] or 0,3)) for k,v in b["config"]["kernels"].items()})
P
