#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_streamed.py tests/test_gpu_c4_full_size.py -m gpu -x -q > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/q_tests.log
