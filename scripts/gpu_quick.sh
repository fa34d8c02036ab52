#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_statevec.py tests/test_gpu_qsplit.py -x -q -m gpu 2>&1 | tail -3
for bb in 4 8 16 32; do echo "QDC_SV_TENSOR_BB=$bb"; QDC_SV_TENSOR_BB=$bb timeout 300 python scripts/probe_variant_sims.py 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['prep_side']['ms'], d['meas_side']['ms'], d['value'])"; done | tee gpurun_out/tensor_bb.log
