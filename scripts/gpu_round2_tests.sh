#!/bin/bash
# whole GPU suite; new files first so that a failure there shows early
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_sampler.py tests/test_gpu_cutqc_pipeline.py -x -q -m gpu 2>&1 | tail -25 | tee gpurun_out/gputests_new.log
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -25 | tee gpurun_out/gputests_all.log
