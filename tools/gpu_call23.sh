#!/bin/bash
# GPU call (development): traceback variants -- one thread per window without / with prefetch (ab/lib?.so), the warp-per-window
# kernel (CUDAPOA_B200_TRACE=warp) -- then the GPU tests on the working tree's library
cd "$(dirname "$0")/.."
VARIANTS="${VARIANTS:-A B C D}" SIZES="" bash tools/ab_variants.sh
echo "== warp-per-window traceback"
CUDAPOA_B200_TRACE=warp python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['kernels']['graph_kernels']
print('warp', 'value %.0f e2e %.0f ms %.1f fill %.1f trace %.1f fuse %.1f sort %.1f final %.1f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['kernels']['poa_fill_kernel']['ms_per_step'], g['poa_trace_kernel_ms'], g['poa_fuse_kernel_ms'], g['poa_sort_kernel_ms'], g['poa_final_kernel_ms']))"
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
