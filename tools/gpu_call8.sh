#!/bin/bash
# GPU call: GPU test suite, bench, end-to-end sweep of the window pipeline, adaptive-banding agreement study.
cd "$(dirname "$0")/.."
TAG=${1:-c8}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    g=d["kernels"]["graph_kernels"]
    print("BENCH value %.0f e2e %.0f ms %.1f fill %.1f graph %.1f (trace %.1f fuse %.1f sort %.1f final %.1f) frac %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["kernels"]["poa_fill_kernel"]["ms_per_step"], g["ms_per_step"], g["poa_trace_kernel_ms"], g["poa_fuse_kernel_ms"], g["poa_sort_kernel_ms"], g["poa_final_kernel_ms"], d["roofline"]["frac"]))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
timeout 900 python tools/pipeline_sweep.py 18944 8,1184,6 8,2368,8 4,2368,8 4,4736,14 2,9472,26 3,6315,20 16,1184,4 > gpurun_out/${TAG}_sweep.log 2>&1
tail -9 gpurun_out/${TAG}_sweep.log
timeout 1500 python tools/adaptive_agreement.py ${N_STRUCT:-2048} ${N_CFG2:-20480} > gpurun_out/${TAG}_adaptive.json 2> gpurun_out/${TAG}_adaptive.err
tail -5 gpurun_out/${TAG}_adaptive.err | cut -c1-700
ls -la gpurun_out | tail -6
