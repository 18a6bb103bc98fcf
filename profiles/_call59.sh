set -x
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 10 --no-e2e > gpurun_out/r4l_bench_n2.json 2> gpurun_out/r4l_bench_n2.err
tail -3 gpurun_out/r4l_bench_n2.err
python - <<'P'
import json
b=json.loads(open("gpurun_out/r4l_bench_n2.json").read().strip().splitlines()[-1])
print(b["value"], b["ms_per_step"], b["config"]["rows_out_per_step"], b.get("nvlink"), b["config"]["phases_ms"])
P
