#!/bin/bash
# GPU call (development): GPU tests, short bench (headline only), one `ncu --set full` capture of a mid-batch fill launch
cd "$(dirname "$0")/.."
TAG=${1:-r2o}
B="python bench.py --no-cpu-baseline --no-other-configs"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 900 $B --steps 5 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    g=d["kernels"]["graph_kernels"]
    print("BENCH value %.0f e2e %.0f ms %.1f fill %.1f graph %.1f (trace %.1f fuse %.1f sort %.1f final %.1f) frac %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["kernels"]["poa_fill_kernel"]["ms_per_step"], g["ms_per_step"], g["poa_trace_kernel_ms"], g["poa_fuse_kernel_ms"], g["poa_sort_kernel_ms"], g["poa_final_kernel_ms"], d["roofline"]["frac"]))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
for k in ${NCU_KERNELS:-fill}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:poa_${k}_kernel -s 14 -c 1 -f -o gpurun_out/${TAG}_${k} $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_${k}.log 2>&1
  echo "ncu $k rc=$?"
done
