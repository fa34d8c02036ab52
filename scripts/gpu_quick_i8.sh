#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_contract_score.py -x -q -m gpu 2>&1 | tail -2
timeout 600 python scripts/time_i8_debug.py 2>&1 | tee gpurun_out/r2_i8_quick.log | tail -12
