#!/bin/bash
# GPU call 1 (round 2): goldens for the new cases from the reference build, the GPU test suite, the reference timed
# with K concurrent batches, DRAM traffic of the reference kernels.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt 2>&1
timeout 900 python tests/golden/make_reference_golden.py cfg4_banded32_real cfg5b_mixed banded_size32 banded_size32_msa > gpurun_out/c1_golden.log 2>&1
echo "golden rc=$?" >> gpurun_out/c1_golden.log
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/c1_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
timeout 600 python tools/reference_concurrent.py --engine ref --ks 1,2,4,8,16 > gpurun_out/c1_ref_concurrent.jsonl 2> gpurun_out/c1_ref_concurrent.err
timeout 300 python tools/reference_concurrent.py --engine native --ks 1,4,8 > gpurun_out/c1_native_concurrent.jsonl 2> gpurun_out/c1_native_concurrent.err
timeout 300 python tools/reference_concurrent.py --engine ref --ks 1 --windows 4736 --reps 1 > gpurun_out/c1_ref_plain.log 2>&1 && \
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:generate -c 6 --csv --log-file gpurun_out/c1_ref_ncu.csv python tools/reference_concurrent.py --engine ref --ks 1 --windows 4736 --reps 1 > gpurun_out/c1_ref_ncu.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err
tail -3 gpurun_out/c1_ref_concurrent.jsonl
cat gpurun_out/c1_bench.json | cut -c1-600
