set -x
B="python bench.py --parts filter --steps 10 --no-e2e --no-cpu-baseline"
timeout 300 $B > gpurun_out/r4c_filter_staged.json 2> gpurun_out/r4c_err.txt
timeout 300 $B --no-staged-filter > gpurun_out/r4c_filter_old.json 2>> gpurun_out/r4c_err.txt
RFGPU_LIB=$PWD/.wip/librfgpu_m5.so timeout 300 $B > gpurun_out/r4c_filter_staged_m5.json 2>> gpurun_out/r4c_err.txt
RFGPU_LIB=$PWD/.wip/librfgpu_m8.so timeout 300 $B > gpurun_out/r4c_filter_staged_m8.json 2>> gpurun_out/r4c_err.txt
tail -3 gpurun_out/r4c_err.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "filter or scan or compatible or minus" 2>&1 | tail -5 > gpurun_out/r4c_tests.txt
cat gpurun_out/r4c_tests.txt
python - <<'P'
import json
for n in ("staged","old","staged_m5","staged_m8"):
    try:
        d=json.loads(open("gpurun_out/r4c_filter_%s.json"%n).read().strip().splitlines()[-1])
        k=d["config"]["kernels"].get("k_filter_compact")
        print(n, d["ms_per_step"], k)
    except Exception as e:
        print(n, "ERR", e)
P
