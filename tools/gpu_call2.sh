#!/bin/bash
# GPU call: GPU test suite, bench (full line), ncu launch list, one full ncu capture of the fill and of the sort kernel.
cd "$(dirname "$0")/.."
TAG=${1:-c2}
B="python bench.py --no-cpu-baseline --no-other-configs"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 ${BENCH_FLAGS:-} > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    g=d["kernels"]["graph_kernels"]
    print("BENCH value %.0f e2e %.0f ms %.1f fill %.1f graph %.1f (trace %.1f fuse %.1f sort %.1f final %.1f) frac %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["kernels"]["poa_fill_kernel"]["ms_per_step"], g["ms_per_step"], g["poa_trace_kernel_ms"], g["poa_fuse_kernel_ms"], g["poa_sort_kernel_ms"], g["poa_final_kernel_ms"], d["roofline"]["frac"]))
    for k,v in (d.get("configs") or {}).items(): print(k, json.dumps(v)[:400])
    print("cpu_baseline", json.dumps(d.get("cpu_baseline"))[:300])
    print("ref", json.dumps(d.get("reference_cudapoa_gpu"))[:300])
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
if [ -z "$NO_NCU" ]; then
timeout 600 $B --steps 1 --warmup 0 > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1300 --csv --log-file gpurun_out/${TAG}_launches.csv $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_fill_kernel -s 15 -c 1 -f -o gpurun_out/${TAG}_fill $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_fill.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_sort_kernel -s 30 -c 1 -f -o gpurun_out/${TAG}_sort $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_sort.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_trace_kernel -s 15 -c 1 -f -o gpurun_out/${TAG}_trace $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_trace.log 2>&1
fi
ls -la gpurun_out | tail -14
