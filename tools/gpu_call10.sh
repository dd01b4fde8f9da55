#!/bin/bash
# development helper (GPU box): GPU tests, sort-class sweep, other shapes under both schedules
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
show() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['kernels']['graph_kernels']
print('$1', 'value %.0f e2e %.0f ms %.1f fill %.1f trace %.1f fuse %.1f sort %.1f final %.1f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['kernels']['poa_fill_kernel']['ms_per_step'], g['poa_trace_kernel_ms'], g['poa_fuse_kernel_ms'], g['poa_sort_kernel_ms'], g['poa_final_kernel_ms']))"; }
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs"
$B 2>/dev/null | show default
for s in 1200 1536; do CUDAPOA_B200_SORT_NODES=$s $B 2>/dev/null | show "sort_nodes=$s"; done
timeout 1200 python tools/other_configs_modes.py cfg4:592 cfg4:2368 cfg5b:2368 cfg5b:9472 2>&1 | grep -v Warning
