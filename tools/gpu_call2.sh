#!/bin/bash
# GPU call: GPU test suite, bench, ncu launch list, one full ncu capture of the fill and of the graph kernel.
cd "$(dirname "$0")/.."
TAG=${1:-c2}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("BENCH value %.0f e2e %.0f ms %.1f fill %.1f graph %.1f frac %.4f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["kernels"]["poa_fill_kernel"]["ms_per_step"], d["kernels"]["poa_graph_kernel"]["ms_per_step"], d["roofline_dpx"]["frac"]))
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
PY
timeout 600 python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_fill_kernel -s 15 -c 1 -f -o gpurun_out/${TAG}_fill python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/${TAG}_ncu_fill.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:poa_graph_kernel -s 30 -c 1 -f -o gpurun_out/${TAG}_graph python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/${TAG}_ncu_graph.log 2>&1
ls -la gpurun_out | tail -12
