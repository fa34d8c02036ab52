#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_streamed.py -m gpu -x -q > gpurun_out/q_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/q_tests.log
timeout 600 python scripts/probe_streamed.py > gpurun_out/streamed_n1.log 2>&1; echo "probe rc=$?"; tail -12 gpurun_out/streamed_n1.log
