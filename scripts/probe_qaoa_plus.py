"""QAOA+ uncut cost (ansatz/qaoa_plus.py) at 26-30 qubits: groups the compiler makes and time per parameter vector."""
import importlib, os, sys, time
import numpy as np, torch, networkx as nx
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import mis
from qdc_b200.ansatz import qaoa_plus
from qdc_b200.circuit import Parameter
from qdc_b200.engine import B200Backend
dev = torch.device("cuda:0")
for n, P in ((26, 1), (28, 1), (30, 1)):
    G = nx.random_regular_graph(3, n, seed=1)
    syms = [Parameter(f"var_{i}") for i in range(2 * P)]
    circ = qaoa_plus.construct_qaoa_plus(P, G, syms)
    be = B200Backend(device=dev, max_state_bytes=40 << 30)
    prog, auto = be.compile(circ, param_order=circ.parameters)
    table = np.random.default_rng(0).uniform(0, 1, (2, 2 * P))
    mis.qaoa_plus_cost(circ, be, G, 2.0, params=table[:1], param_order=circ.parameters)
    torch.cuda.synchronize(); t = time.time()
    out = mis.qaoa_plus_cost(circ, be, G, 2.0, params=table, param_order=circ.parameters)
    torch.cuda.synchronize(); dt = (time.time() - t) / 2
    print(f"QAOA+ n={n} P={P}: groups {prog.n_groups}, ops {prog.n_ops}, {dt * 1e3:.2f} ms per parameter vector, cost {out[0]:.6f}", flush=True)
