#!/bin/bash
# GPU call: the headline shape at other batch sizes (windows per step), resident and end to end
cd "$(dirname "$0")/.."
show() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); g=d['kernels']['graph_kernels']
print('$1', 'value %.0f e2e %.0f ms %.1f fill %.1f trace %.1f fuse %.1f sort %.1f final %.1f frac %.4f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['kernels']['poa_fill_kernel']['ms_per_step'], g['poa_trace_kernel_ms'], g['poa_fuse_kernel_ms'], g['poa_sort_kernel_ms'], g['poa_final_kernel_ms'], d['roofline']['frac']))"; }
for n in ${SIZES:-4736 9472 37888}; do
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs --windows $n --max-mem-gb 110 2>gpurun_out/sweep.err | show "windows=$n"
  tail -2 gpurun_out/sweep.err | cut -c1-200
done
