#!/bin/bash
# GPU call (2 GPUs): the driver's round-end commands at small step counts -- reference arm at N=1, both arms under torchrun at N=2
cd "$(dirname "$0")/.."
TAG=${1:-r2v}
mkdir -p gpurun_out
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc=$?"
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_2gpu.json 2> gpurun_out/${TAG}_bench_2gpu.err; echo "2gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_2gpu.json 2> gpurun_out/${TAG}_bench_reference_2gpu.err; echo "ref 2gpu rc=$?"
python - <<PY
import json
for f in ("bench_2gpu","bench_reference","bench_reference_2gpu"):
    try:
        d=json.loads(open("gpurun_out/${TAG}_%s.json"%f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ("value","n_gpus","ms_per_step","impl")}, "e2e", (d.get("e2e") or {}).get("value"), "frac", (d.get("roofline") or {}).get("frac"))
        for k,v in (d.get("configs") or {}).items(): print("   ", k, "value %.0f e2e %.0f ms %.1f" % (v["value"], v["e2e"], v["ms_per_step"]))
    except Exception as e:
        print(f, "parse failed", e); print(open("gpurun_out/${TAG}_%s.err"%f).read()[-1500:])
PY
