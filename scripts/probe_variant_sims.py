"""bench.variant_sims_leg alone (for an ncu launch list of the statevector kernels at the C4 fragment shapes)."""
import importlib, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
import bench
out = bench.variant_sims_leg(torch, torch.device("cuda:0"), cpu=False)
print(json.dumps({k: out[k] for k in ("prep_side", "meas_side", "value", "ms_total")}))
