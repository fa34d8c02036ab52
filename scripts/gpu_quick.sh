#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
  scripts/probe_e2e_phases.py > gpurun_out/e2e_phases_n$N.log 2>&1; echo "rc=$?"
tail -22 gpurun_out/e2e_phases_n$N.log
