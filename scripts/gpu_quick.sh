#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_statevec.py tests/test_gpu_configs.py tests/test_gpu_solver.py -x -q -m gpu 2>&1 | tail -12
timeout 600 python bench.py --workload c5 --param-sets 8 --steps 2 > gpurun_out/bench_c5_fused.json 2> gpurun_out/bench_c5_fused.err; echo "c5 rc=$?"; tail -c 400 gpurun_out/bench_c5_fused.err
tail -n 1 gpurun_out/bench_c5_fused.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline'])"
