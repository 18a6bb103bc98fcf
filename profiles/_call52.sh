set -x
B="python bench.py --steps 10 --no-e2e --no-cpu-baseline"
timeout 300 $B --parts filter > gpurun_out/r4e_filter.json 2> gpurun_out/r4e_err.txt
tail -2 gpurun_out/r4e_err.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "filter or scan or compatible or wide" 2>&1 | tail -15 > gpurun_out/r4e_tests.txt
cat gpurun_out/r4e_tests.txt
python - <<'P'
import json
d=json.loads(open("gpurun_out/r4e_filter.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["config"]["kernels"].get("k_filter_compact"), d["config"].get("rows_out_per_step"))
P
