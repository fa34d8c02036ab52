#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_cutqc.py tests/test_gpu_cutqc_pipeline.py -x -q -m gpu 2>&1 | tail -4
timeout 600 python scripts/time_cutqc.py 2>&1 | grep -E "measure|merge" | tee gpurun_out/time_marginal.log
