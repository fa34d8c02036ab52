#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cutqc.py tests/test_gpu_cutqc_pipeline.py -m gpu -x -q > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/q_tests.log
timeout 600 python scripts/time_cutqc.py > gpurun_out/time_cutqc.log 2>&1; echo "rc=$?"; grep -E "measure|merge|collapse" gpurun_out/time_cutqc.log | tail -12
QDC_MARG_DIRECT=0 timeout 600 python scripts/time_cutqc.py 2>&1 | grep -E "measure|merge" | tail -8
