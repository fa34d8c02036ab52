#!/bin/bash
mkdir -p gpurun_out
touch dqva-and-circuit-cutting_b200/csrc/contract_i8.cu
make -C dqva-and-circuit-cutting_b200/csrc EXTRA=-DQDC_I8_TRACE 2>&1 | tail -1
timeout 600 python scripts/trace_i8.py ${1:-5} > gpurun_out/i8_trace.log 2>&1; echo "rc=$?"
awk '/pair=1/{f=1} f' gpurun_out/i8_trace.log | grep -E "tile 20[12] +epi" | head -60
