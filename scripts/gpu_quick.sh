#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/probe_marginal.py 2>&1 | tail -2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:marginal_kernel -o gpurun_out/r2_marginal python scripts/probe_marginal.py > gpurun_out/ncu_marg.log 2>&1; echo rc=$?
