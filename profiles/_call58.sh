set -x
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "search_mode or lazy or lookup" 2>&1 | tail -5 > gpurun_out/r4k_lookup_tests.txt
cat gpurun_out/r4k_lookup_tests.txt
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r4k_bench.json 2> gpurun_out/r4k_bench.err
python - <<'P'
import json
b=json.loads(open("gpurun_out/r4k_bench.json").read().strip().splitlines()[-1])
print(b["value"], b["ms_per_step"], b["e2e"]["value"], b["config"]["rows_out_per_step"], b["config"]["phases_ms"], {k:(v["launches"], round(v["ms_per_launch"]*1e3,1), round(v["frac"] or 0,3)) for k,v in b["config"]["kernels"].items()})
P
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r4k_gpu_tests.txt
tail -3 gpurun_out/r4k_gpu_tests.txt
