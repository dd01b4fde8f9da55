#!/bin/bash
# GPU call (8 GPUs): the bench line at N = 8 under torchrun
cd "$(dirname "$0")/.."
TAG=${1:-r2z}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 3 --warmup 3 --no-other-configs > gpurun_out/${TAG}_bench_8gpu.json 2> gpurun_out/${TAG}_bench_8gpu.err; echo "8gpu rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench_8gpu.json").read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("value","n_gpus","ms_per_step")}, "e2e", d["e2e"]["value"], "frac", d["roofline"]["frac"], "cores", (d.get("cpu_baseline") or {}).get("cores"))
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/${TAG}_bench_8gpu.err").read()[-1500:])
PY
