#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python scripts/time_configs.py 2>&1 | tail -25 | tee gpurun_out/time_configs.log
