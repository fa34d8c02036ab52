#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 \
  scripts/probe_streamed.py > gpurun_out/streamed_n$N.log 2>&1; echo "probe rc=$?"; grep -E "^N=|Error|error" gpurun_out/streamed_n$N.log | tail -12
