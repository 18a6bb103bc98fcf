set -x
B="python bench.py --steps 10 --no-e2e --no-cpu-baseline"
for v in oldfilter filterA; do
  RFGPU_LIB=$PWD/.wip/librfgpu_$v.so timeout 300 $B > gpurun_out/r4g_full_$v.json 2>> gpurun_out/r4g_err.txt
done
timeout 300 $B > gpurun_out/r4g_full_filterB.json 2>> gpurun_out/r4g_err.txt
timeout 300 $B --parts filter > gpurun_out/r4g_part_filterB.json 2>> gpurun_out/r4g_err.txt
tail -2 gpurun_out/r4g_err.txt
python - <<'P'
import json
for n in ("full_oldfilter","full_filterA","full_filterB","part_filterB"):
    try:
        d=json.loads(open("gpurun_out/r4g_%s.json"%n).read().strip().splitlines()[-1])
        k=d["config"]["kernels"]["k_filter_compact"]
        print(n, round(d["ms_per_step"],4), round(k["ms_per_launch"]*1e3,1), round(k["frac"],3))
    except Exception as e:
        print(n,"ERR",e)
P
