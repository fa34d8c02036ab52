#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_statevec.py tests/test_gpu_qsplit.py tests/test_gpu_configs.py tests/test_gpu_sampler.py -x -q -m gpu 2>&1 | tail -5
timeout 600 python scripts/time_uncut.py 29 1 2>&1 | tail -5 | tee gpurun_out/sv_bulk_n29.log
timeout 600 python scripts/time_uncut.py 25 2 2>&1 | tail -5 | tee gpurun_out/sv_bulk_n25.log
timeout 600 python bench.py --workload c5 --param-sets 8 --steps 2 > gpurun_out/bench_c5_bulk.json 2> gpurun_out/bench_c5_bulk.err; echo "c5 rc=$?"; tail -c 600 gpurun_out/bench_c5_bulk.json
