#!/bin/bash
# final N = 1 evidence: bench line, launch list of a short bench run, one full capture of the contraction kernel
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
timeout 900 python bench.py --workload c4-circuit --param-sets 4 --steps 3 > gpurun_out/bench_c4_circuit.json 2> gpurun_out/bench_c4_circuit.err; echo "c4-circuit rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 8 -c 1 -o gpurun_out/r2_i8_pair_final \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
tail -n 1 gpurun_out/bench_default.json | python -c "
import sys,json; d=json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['parity_check']['pass'], d['clocks'])
print('c4_circuit', d['c4_circuit']['ms_per_evaluation'], 'c5', d['c5_sample']['ms_per_state'], 'variant', d['variant_sims']['value'])"
tail -n 1 gpurun_out/bench_c4_circuit.json | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('c4-circuit', d['value'], d['ms_per_step'], d['result']['ms_per_evaluation'])"
