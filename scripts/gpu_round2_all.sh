#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -6 | tee gpurun_out/gputests_all.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -n 1 gpurun_out/bench_default.json | python -c "
import sys,json; d=json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['parity_check']['pass'], d['clocks'], d['gpu_launches'])
print({k:(v.get('ms_per_step'), v.get('n_candidates')) for k,v in d['other_methods']['peaked_operands'].items() if isinstance(v, dict)}, d['other_methods']['peaked_operands'].get('same_top'))
print('c4_circuit', d['c4_circuit']['ms_per_evaluation'], 'c5', d['c5_sample']['ms_per_state'], 'variant', d['variant_sims']['value'], d['variant_sims']['roofline']['frac'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -n 1 | cut -c1-300
