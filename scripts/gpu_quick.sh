#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_qsplit.py tests/test_gpu_configs.py -x -q -m gpu 2>&1 | tail -5
timeout 600 python scripts/time_c4_circuit.py > gpurun_out/c4_stages_plain.log 2>&1; tail -7 gpurun_out/c4_stages_plain.log
