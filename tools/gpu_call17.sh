#!/bin/bash
# GPU call (development): GPU tests on the working tree's library, then the kernel variants in ab/lib?.so alternately
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
VARIANTS="${VARIANTS:-A B C}" SIZES="${SIZES:-}" bash tools/ab_variants.sh
