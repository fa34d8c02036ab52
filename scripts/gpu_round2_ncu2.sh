#!/bin/bash
# full capture (with source) of the int8 contraction kernel at TN = 128 (default) and TN = 64
set -u
mkdir -p gpurun_out
CMD="python scripts/time_i8_debug.py"
$CMD > gpurun_out/r2_plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 4 -c 1 -f -o gpurun_out/r2_i8_full_tn128 $CMD > gpurun_out/r2_ncu4.log 2>&1
echo "ncu tn128 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 12 -c 1 -f -o gpurun_out/r2_i8_full_tn64 $CMD > gpurun_out/r2_ncu5.log 2>&1
echo "ncu tn64 rc=$?"
