#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cutqc_pipeline.py tests/test_gpu_cutqc_stages.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python scripts/time_cutqc.py 2>&1 | tee gpurun_out/time_cutqc.log | head -12
