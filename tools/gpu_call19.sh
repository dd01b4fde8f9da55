#!/bin/bash
# GPU call (development): GPU tests, end-to-end sweep of the double-buffered window pipeline, one traced run
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/pipeline_sweep.py 18944 ${SWEEP:-8,1184,12 4,1184,12 6,1184,12 4,2368,20 8,592,6 2,2368,20 3,1579,16 4,1579,16} 2>&1 | grep -v Warning
timeout 600 python tools/pipeline_trace.py 18944 ${TRACE:-4,1184,12} 2>&1 | grep -v Warning
