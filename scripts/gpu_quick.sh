#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/probe_cut_cost.py 2>&1 | tail -45 | tee gpurun_out/cut_cost_profile.log
