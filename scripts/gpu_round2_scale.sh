#!/bin/bash
# bench.py under torchrun at N ranks (N = number of visible GPUs), plus the reference arm on rank 0
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_scale_n$N.json 2> gpurun_out/r2_scale_n$N.err
echo "bench N=$N rc=$?"
tail -3 gpurun_out/r2_scale_n$N.err
python - <<P
import json
try:
    d=json.loads(open("gpurun_out/r2_scale_n$N.json").read().strip().split("\n")[-1]); r=d["roofline"]
    print('N=$N value',round(d['value']/1e9,1),'Gprobs/s ms_per_step',round(d['ms_per_step'],3),'kernel_ms',round(r['kernel_ms'],3),'e2e ms',round(d['e2e']['ms_per_step'],3),'e2e value',round(d['e2e']['value']/1e9,1),'h2d',d['e2e']['h2d_bytes_per_step'],'launches',d['gpu_launches'],'cost',d['result']['avg_cost'])
except Exception as e: print('parse failed',e)
P
