#!/bin/bash
# GPU call: GPU tests (incl. packed full alignment), cfg2 in full-alignment mode under both schedules, the reference build
# in full-alignment mode driven with K concurrent batches
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/other_configs_modes.py cfg2_full:2368 cfg2_full:9472 cfg3:4736 cfg5b:2368 2>&1 | grep -v Warning
timeout 900 python tools/reference_concurrent.py --engine ref --config cfg2 --full --windows 8192 --ks 4,8,16 --reps 1 2>&1 | tail -4
