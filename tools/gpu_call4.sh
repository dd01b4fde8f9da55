#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python tools/debug_mode2.py 4 > gpurun_out/c4_debug.log 2>&1
cat gpurun_out/c4_debug.log | tail -30
timeout 900 compute-sanitizer --tool racecheck --racecheck-report all python tools/debug_mode2.py 1 banded16_short_reads > gpurun_out/c4_racecheck.log 2>&1
grep -E "RACECHECK SUMMARY|Race reported|ERROR SUMMARY" gpurun_out/c4_racecheck.log | head -20
