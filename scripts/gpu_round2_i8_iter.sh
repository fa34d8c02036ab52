#!/bin/bash
# one iteration on the int8 contraction: targeted tests, then the bench at both tile widths
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_contract_score.py tests/test_gpu_cutqc.py -x -q -m gpu > gpurun_out/r2_iter_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r2_iter_tests.log
for tn in 128 96 64; do
  QDC_I8_TN=$tn timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_iter_tn$tn.json 2> gpurun_out/r2_iter_tn$tn.err
  echo "bench tn=$tn rc=$?"
  python - <<P
import json
d=json.load(open('gpurun_out/r2_iter_tn$tn.json')); r=d['roofline']
print('tn=$tn ms_per_step',round(d['ms_per_step'],3),'kernel_ms',round(r['kernel_ms'],3),'e2e',round(d['e2e']['ms_per_step'],3),'TOP/s',round(r['achieved'],1),'clk',d['clocks']['sm_mhz'],d['clocks']['reasons'],'cost',d['result']['avg_cost'],{k:round(v['ms_per_step'],2) for k,v in d['other_methods'].items()})
P
done
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2_noalloc_alone.log 2>&1
echo "no-alloc alone rc=$?"; tail -2 gpurun_out/r2_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_t_all2.log 2>&1
echo "all rc=$?"; tail -3 gpurun_out/r2_t_all2.log
