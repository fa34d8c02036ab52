#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_statevec.py tests/test_gpu_sampler.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python scripts/time_uncut.py 29 1 2>&1 | tail -5 | tee gpurun_out/sv_bulk_n29.log
timeout 600 python scripts/time_uncut.py 25 2 2>&1 | tail -5 | tee gpurun_out/sv_bulk_n25.log
timeout 300 python scripts/probe_variant_sims.py 2>&1 | tail -1 | cut -c1-400
