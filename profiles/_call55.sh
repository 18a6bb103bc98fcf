set -x
for i in 1 2; do
  for v in oldfilter filterA; do
    RFGPU_LIB=$PWD/.wip/librfgpu_$v.so timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r4h_default_${v}_$i.json 2>> gpurun_out/r4h_err.txt
  done
done
python - <<'P'
import json
for i in (1,2):
  for n in ("oldfilter","filterA"):
    try:
        d=json.loads(open("gpurun_out/r4h_default_%s_%d.json"%(n,i)).read().strip().splitlines()[-1])
        k=d["config"]["kernels"]["k_filter_compact"]
        print(n, i, round(d["ms_per_step"],4), round(k["ms_per_launch"]*1e3,1), round(k["frac"],3), round(d["e2e"]["ms_per_step"],3))
    except Exception as e:
        print(n,"ERR",e)
P
