#!/bin/bash
# GPU call: evidence at one commit -- smoke, GPU tests, full bench line, ncu launch list of one bench pass,
# one `ncu --set full` capture each of a mid-batch poa_fill_kernel / poa_trace_kernel / poa_sort_kernel launch
cd "$(dirname "$0")/.."
TAG=${1:-r2n}
B="python bench.py --no-cpu-baseline --no-other-configs"
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 1200 python bench.py ${BENCH_FLAGS:-} > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    g=d["kernels"]["graph_kernels"]
    print("BENCH value %.0f e2e %.0f ms %.1f fill %.1f graph %.1f (trace %.1f fuse %.1f sort %.1f final %.1f) frac %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["kernels"]["poa_fill_kernel"]["ms_per_step"], g["ms_per_step"], g["poa_trace_kernel_ms"], g["poa_fuse_kernel_ms"], g["poa_sort_kernel_ms"], g["poa_final_kernel_ms"], d["roofline"]["frac"]))
    for k,v in (d.get("configs") or {}).items(): print(k, "value %.0f e2e %.0f ms %.1f" % (v["value"], v["e2e"], v["ms_per_step"]), v.get("phase_ms_per_step"))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
if [ -z "$NO_NCU" ]; then
timeout 600 $B --steps 1 --warmup 0 > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -c 900 --csv --log-file gpurun_out/${TAG}_launches.csv $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_launches.log 2>&1
for k in ${NCU_KERNELS:-fill trace_lanes sort}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:poa_${k}_kernel -s 14 -c 1 -f -o gpurun_out/${TAG}_${k} $B --steps 1 --warmup 0 > gpurun_out/${TAG}_ncu_${k}.log 2>&1
  echo "ncu $k rc=$?"
done
fi
ls -la gpurun_out | tail -12
# compute-sanitizer on the smoke batch (8 windows, both schedules): memcheck and racecheck logs for profiles/
if [ -n "$SANITIZER" ]; then
  for tool in memcheck racecheck; do
    timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/${TAG}_sanitizer_${tool}.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_sanitizer_${tool}.out 2>&1
    echo "compute-sanitizer $tool rc=$?"; tail -3 gpurun_out/${TAG}_sanitizer_${tool}.log
  done
fi
# parity at scale against the live reference build (25 600 distinct windows, seven cases, both schedules)
if [ -n "$PARITY_AT_SCALE" ]; then
  timeout 1500 python tools/parity_at_scale.py gpurun_out/${TAG}_parity_at_scale.json 2>&1 | grep -v Warning | tail -12
fi
