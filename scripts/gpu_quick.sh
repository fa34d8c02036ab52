#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sv_meas_tensor -s 2 -c 2 -o gpurun_out/r2_meas_tensor python scripts/probe_variant_sims.py > gpurun_out/ncu_mt.log 2>&1; echo rc=$?
