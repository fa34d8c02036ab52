#!/bin/bash
# default bench line (with the circuit-level and C5 sub-objects) + the two extra workloads
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "default rc=$?"
tail -c 3000 gpurun_out/bench_default.err
timeout 600 python bench.py --workload c4-circuit --param-sets 4 --steps 3 > gpurun_out/bench_c4_circuit.json 2> gpurun_out/bench_c4_circuit.err; echo "c4-circuit rc=$?"
tail -c 2000 gpurun_out/bench_c4_circuit.err
timeout 900 python bench.py --workload c5 --param-sets 8 --steps 1 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"
tail -c 2000 gpurun_out/bench_c5.err
python - <<'P'
import json
for f in ("bench_default","bench_c4_circuit","bench_c5"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d.get("value"), d.get("ms_per_step"), {k:d[k] for k in ("c4_circuit","c5_sample") if k in d})
        if "variant_sims" in d: print(" vs", d["variant_sims"].get("value"), d["variant_sims"].get("cpu_baseline"))
        if "other_methods" in d: print(" peaked", d["other_methods"].get("peaked_operands"))
    except Exception as e: print(f, "unreadable", e)
P
