#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -6 | tee gpurun_out/gputests_all.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -n 1 gpurun_out/bench_default.json | python -c "
import sys,json; d=json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['e2e'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline']['frac_of_measured_int8_issue_peak'], d['parity_check']['pass'], d['clocks'])
print(d['other_methods']['peaked_operands'])
print(d['c4_circuit']['ms_per_evaluation'], d['c5_sample']['ms_per_state'])"
