#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_statevec.py tests/test_gpu_configs.py -m gpu -x -q > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?" 
tail -3 gpurun_out/q_tests.log
python scripts/probe_qaoa_plus.py > gpurun_out/qaoa_plus_probe3.log 2>&1; tail -8 gpurun_out/qaoa_plus_probe3.log
python scripts/time_uncut.py 29 1 > gpurun_out/uncut29.log 2>&1; tail -4 gpurun_out/uncut29.log
