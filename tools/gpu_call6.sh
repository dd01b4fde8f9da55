#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in 2 3 4 6 10; do
  python tools/debug_dump.py banded16_short_reads 0 $n gpurun_out/gpu_dump_$n.npz
done
