#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/time_c4_circuit.py 2>&1 | tail -8 | tee gpurun_out/c4_circuit_stages.log
