#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for f in 0 1 2 4 8 15; do
  echo "== flags $f" >> gpurun_out/c5_debug.log
  CUDAPOA_B200_DEBUG_FLAGS=$f timeout 300 python tools/debug_mode2.py 4 banded16_short_reads cfg2_banded16_consensus 2>&1 | grep "mode 2" >> gpurun_out/c5_debug.log
done
cat gpurun_out/c5_debug.log
