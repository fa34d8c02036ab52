#!/bin/bash
mkdir -p gpurun_out
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 10 -c 1 -o gpurun_out/r2_i8_pair_final \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
