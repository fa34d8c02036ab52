#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_qsplit.py tests/test_gpu_contract_score.py -x -q -m gpu > gpurun_out/r2_ml_tests.log 2>&1
echo "ml+contract tests rc=$?"; tail -15 gpurun_out/r2_ml_tests.log
timeout 600 python scripts/time_i8_debug.py > gpurun_out/r2_i8_debug.log 2>&1
echo "debug rc=$?"; cat gpurun_out/r2_i8_debug.log | tail -20
CMD2="python bench.py --steps 1 --warmup 3 --slices 5 --no-cpu-baseline"
$CMD2 > gpurun_out/r2_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 3 -c 1 -f -o gpurun_out/r2_i8_full_lean $CMD2 > gpurun_out/r2_ncu3.log 2>&1
echo "ncu rc=$?"
