#!/bin/bash
# launch list of the default bench + one full capture (with source) of the int8 contraction kernel
set -u
mkdir -p gpurun_out
CMD1="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD1 > gpurun_out/r2_plain1.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD1 > gpurun_out/r2_ncu1.log 2>&1
echo "launch list rc=$?"
CMD2="python bench.py --steps 1 --warmup 3 --slices 5 --no-cpu-baseline"
$CMD2 > gpurun_out/r2_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 3 -c 1 -f -o gpurun_out/r2_i8_full $CMD2 > gpurun_out/r2_ncu2.log 2>&1
echo "full rc=$?"
ls -la gpurun_out/ | tail -8
