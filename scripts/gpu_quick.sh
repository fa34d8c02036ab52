#!/bin/bash
mkdir -p gpurun_out
for g in 0 6 12 24; do echo "grid per SM $g"; QDC_MARG_DIRECT_GRID=$g timeout 600 python scripts/time_cutqc.py 2>&1 | grep -E "measure" | tail -3; done
