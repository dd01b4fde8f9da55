#!/bin/bash
# development helper (GPU box): time ab/libA.so and ab/libB.so alternately on the same device
cd "$(dirname "$0")/.."
L=claragenomicsanalysis_b200/lib/libcudapoa_b200.so
cp $L /tmp/keep.so
for round in 1 2; do
for v in A B; do
  cp ab/lib$v.so $L
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', round(d['value']), round(d['e2e']['value']))"
done
done
cp /tmp/keep.so $L
