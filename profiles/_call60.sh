set -x
timeout 60 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/r4m_bench.json 2> gpurun_out/r4m_bench.err
python - <<'P'
import json
b=json.loads(open("gpurun_out/r4m_bench.json").read().strip().splitlines()[-1])
print(b["value"], b["ms_per_step"], b["config"]["rows_out_per_step"], b["config"]["phases_ms"], {k:(v["launches"], round(v["ms_per_launch"]*1e3,1), round(v["frac"] or 0,3)) for k,v in b["config"]["kernels"].items()})
P
timeout 150 python -m pytest tests/test_gpu_parity.py tests/test_gpu_synth.py tests/test_gpu_zy_plan_shapes.py -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r4m_tests.txt
tail -3 gpurun_out/r4m_tests.txt
