#!/bin/bash
# development helper (GPU box): bench the kernel variants in ab/lib?.so alternately, then the default one at other batch sizes
cd "$(dirname "$0")/.."
L=claragenomicsanalysis_b200/lib/libcudapoa_b200.so
cp $L /tmp/keep.so
show() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['kernels']['graph_kernels']
print('$1', 'value %.0f e2e %.0f ms %.1f fill %.1f trace %.1f fuse %.1f sort %.1f final %.1f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['kernels']['poa_fill_kernel']['ms_per_step'], g['poa_trace_kernel_ms'], g['poa_fuse_kernel_ms'], g['poa_sort_kernel_ms'], g['poa_final_kernel_ms']))"; }
for round in 1 2; do
for v in ${VARIANTS:-A B C}; do
  cp ab/lib$v.so $L
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs 2>/dev/null | show $v
done
done
cp /tmp/keep.so $L
for n in ${SIZES-17760 35520}; do
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs --windows $n --max-mem-gb 120 2>/dev/null | show "windows=$n"
done
