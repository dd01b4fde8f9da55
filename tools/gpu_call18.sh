#!/bin/bash
# GPU call (development): traced runs of the window pipeline, schedule crossover of lone batches at the current kernels
cd "$(dirname "$0")/.."
timeout 900 python tools/pipeline_trace.py 18944 8,1184,6 4,2368,8 2>&1 | grep -v Warning
echo "== schedule crossover (cfg 2, lone batch)"
timeout 600 python tools/schedule_crossover.py 592 1184 1776 2368 3552 2>&1 | tail -6
echo "== other shapes, both schedules"
timeout 900 python tools/other_configs_modes.py cfg5b:2368 cfg3:2368 2>&1 | grep -v Warning
