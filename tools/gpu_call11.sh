#!/bin/bash
# development helper (GPU box): cfg 4 under both schedules with the wide sort class on/off, end-to-end pipeline experiments
cd "$(dirname "$0")/.."
for wn in 10240 3456 16384; do
  echo "== SORT_WIDE_NODES=$wn"
  CUDAPOA_B200_SORT_WIDE_NODES=$wn timeout 900 python tools/other_configs_modes.py cfg4:592 cfg4:2368 2>&1 | grep "mode 2"
done
echo "== pipeline, default connections"
timeout 600 python tools/pipeline_sweep.py 37888 8,2368,8 12,2368,6 16,1184,4
echo "== pipeline, CUDA_DEVICE_MAX_CONNECTIONS=32"
CUDA_DEVICE_MAX_CONNECTIONS=32 timeout 600 python tools/pipeline_sweep.py 37888 8,2368,8 12,2368,6 16,1184,4 24,1184,3
