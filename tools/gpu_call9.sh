#!/bin/bash
# development helper (GPU box): resident-blocks sweep of the graph phases, then the other shapes under both schedules
cd "$(dirname "$0")/.."
show() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['kernels']['graph_kernels']
print('$1', 'value %.0f e2e %.0f ms %.1f fill %.1f trace %.1f fuse %.1f sort %.1f final %.1f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['kernels']['poa_fill_kernel']['ms_per_step'], g['poa_trace_kernel_ms'], g['poa_fuse_kernel_ms'], g['poa_sort_kernel_ms'], g['poa_final_kernel_ms']))"; }
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs"
$B 2>/dev/null | show default
for s in 6 5 4 3; do CUDAPOA_B200_BLOCKS_SORT=$s $B 2>/dev/null | show "sort=$s"; done
for t in 5 4; do CUDAPOA_B200_BLOCKS_TRACE=$t $B 2>/dev/null | show "trace=$t"; done
for f in 6 4; do CUDAPOA_B200_BLOCKS_FUSE=$f $B 2>/dev/null | show "fuse=$f"; done
CUDAPOA_B200_BLOCKS_FILL=5 $B 2>/dev/null | show "fill=5"
timeout 1200 python tools/other_configs_modes.py cfg4:592 cfg4:2368 cfg5b:2368 cfg5b:9472 cfg3:4736 cfg3:18944 2>&1 | grep -v Warning
