set -x
B="python bench.py --steps 10 --no-e2e --no-cpu-baseline"
RFGPU_LIB=$PWD/.wip/librfgpu_nolb.so timeout 300 $B --parts filter > gpurun_out/r4d_filter_nolookback.json 2> gpurun_out/r4d_err.txt
RFGPU_LIB=$PWD/.wip/librfgpu_nocopy.so timeout 300 $B --parts filter > gpurun_out/r4d_filter_nocopyout.json 2>> gpurun_out/r4d_err.txt
timeout 300 $B --parts star > gpurun_out/r4d_star.json 2>> gpurun_out/r4d_err.txt
RFGPU_LIB=$PWD/.wip/librfgpu_mjnolb.so timeout 300 $B --parts star > gpurun_out/r4d_star_nolookback.json 2>> gpurun_out/r4d_err.txt
tail -3 gpurun_out/r4d_err.txt
python - <<'P'
import json
for n,k in (("filter_nolookback","k_filter_compact"),("filter_nocopyout","k_filter_compact"),("star","k_merge_join"),("star_nolookback","k_merge_join")):
    try:
        d=json.loads(open("gpurun_out/r4d_%s.json"%n).read().strip().splitlines()[-1])
        print(n, d["ms_per_step"], d["config"]["kernels"].get(k))
    except Exception as e:
        print(n, "ERR", e)
P
