set -x
B="python bench.py --steps 10 --no-e2e --no-cpu-baseline"
timeout 300 $B > gpurun_out/r4j_full_shipped.json 2>> gpurun_out/r4j_err.txt
for v in t128 m6 t128m12; do
  RFGPU_LIB=$PWD/.wip/librfgpu_$v.so timeout 300 $B > gpurun_out/r4j_full_$v.json 2>> gpurun_out/r4j_err.txt
done
python - <<'P'
import json
for n in ("shipped","t128","m6","t128m12"):
    try:
        d=json.loads(open("gpurun_out/r4j_full_%s.json"%n).read().strip().splitlines()[-1])
        k=d["config"]["kernels"]["k_filter_compact"]
        print(n, round(d["ms_per_step"],4), round(k["ms_per_launch"]*1e3,1), round(k["frac"],3))
    except Exception as e:
        print(n,"ERR",e)
P
