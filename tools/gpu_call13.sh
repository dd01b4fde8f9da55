#!/bin/bash
# development helper (GPU box): does capping the fill's resident blocks let other streams' graph phases co-run?
cd "$(dirname "$0")/.."
for f in 6 5 4 3; do
  echo "== BLOCKS_FILL=$f"
  CUDAPOA_B200_BLOCKS_FILL=$f timeout 600 python tools/pipeline_sweep.py 37888 2,9472,26 3,6315,18 8,2368,8 2>&1 | grep windows/s
done
echo "== BLOCKS_FILL=4 BLOCKS_SORT=4 BLOCKS_TRACE=4"
CUDAPOA_B200_BLOCKS_FILL=4 CUDAPOA_B200_BLOCKS_SORT=4 CUDAPOA_B200_BLOCKS_TRACE=4 timeout 600 python tools/pipeline_sweep.py 37888 2,9472,26 3,6315,18 8,2368,8 2>&1 | grep windows/s
echo "== schedule crossover (cfg 2, lone batch): windows, single-launch ms, phase-split ms"
timeout 600 python tools/schedule_crossover.py 592 1184 1776 2368 3552 4736 2>&1 | tail -8
