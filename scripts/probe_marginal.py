"""One measure (d = 20, 144 instances, f32 in) and one merge (eff = 22, half merged) call of the tiled marginal kernel, for ncu."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200.cutqc import kernels as ck
dev = torch.device("cuda:0")
d, n_inst = 20, 144
raw = torch.rand(n_inst, 1 << d, dtype=torch.float32).to(dev)
keep = (1 << d) - 1 - 0b11
for _ in range(3):
    ck.measure(raw, keep, [i % 4 for i in range(n_inst)])
eff = 22
vec = torch.rand(64, 1 << eff, dtype=torch.float64).to(dev)
states = [-1 if q % 2 == 0 else -2 for q in range(eff)]      # half active, half merged (as time_cutqc.py)
for _ in range(3):
    try:
        ck.merge(vec, states)
    except Exception as exc:
        print("merge failed:", exc); break
torch.cuda.synchronize()
print("done")
