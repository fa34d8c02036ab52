#!/bin/bash
# first GPU pass of round 2: targeted int8 tests, then the whole GPU suite, then a short bench
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_contract_score.py -x -q -m gpu > gpurun_out/r2_t_contract.log 2>&1
echo "contract rc=$?" | tee -a gpurun_out/r2_rc.txt
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/r2_t_all.log 2>&1
echo "all rc=$?" | tee -a gpurun_out/r2_rc.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_first.json 2> gpurun_out/r2_bench_first.err
echo "bench rc=$?" | tee -a gpurun_out/r2_rc.txt
tail -5 gpurun_out/r2_t_contract.log; tail -5 gpurun_out/r2_t_all.log; head -c 1500 gpurun_out/r2_bench_first.json
