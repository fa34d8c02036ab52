"""Ad-hoc device timing of K5 (fused contraction + MIS epilogue) and K8 (reduction)."""
import importlib, os, sys, time
import numpy as np, torch, networkx as nx
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import recombine as rc

def ev_time(fn, reps=3, warm=1):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e-3)
    return min(ts), ts

dev = torch.device("cuda:0")
rb = int(sys.argv[1]) if len(sys.argv) > 1 else 15
cb = int(sys.argv[2]) if len(sys.argv) > 2 else 17
kdim = 256
n = rb + cb
g = torch.Generator(device="cpu").manual_seed(0)
A = torch.rand(1 << rb, kdim, dtype=torch.float64, generator=g).to(dev)
Bt = torch.rand(1 << cb, kdim, dtype=torch.float64, generator=g).to(dev)
G = nx.random_regular_graph(3, n, seed=1)
adj = [sum(1 << j for j in G.neighbors(i)) for i in range(n)]
tabs = rc.build_tables(rb, cb, list(range(n)), adj, (1 << n) - 1, device=dev)
outs = float(1 << n)
t, ts = ev_time(lambda: rc.contract_and_score(A, Bt, tabs, k=1))
print(f"K5 fused   rb={rb} cb={cb}: {t*1e3:.2f} ms  {outs/t:.3e} probs/s  {2*outs*kdim/t/1e12:.2f} TFLOP/s", ts)
if n <= 33:
    t, ts = ev_time(lambda: rc.contract_and_score(A, Bt, tabs, k=1, store="united", out_dtype=torch.float32))
    print(f"K5 store32 : {t*1e3:.2f} ms  {outs/t:.3e} probs/s  write {outs*4/t/1e9:.1f} GB/s", ts)
if n <= 32:
    res = rc.contract_and_score(A, Bt, tabs, k=1, store="united")
    t, ts = ev_time(lambda: rc.contract_and_score(A, Bt, tabs, k=1, store="united", out=res.P))
    print(f"K5 store64 : {t*1e3:.2f} ms  {outs/t:.3e} probs/s  write {outs*8/t/1e9:.1f} GB/s", ts)
    P = res.P
    t, ts = ev_time(lambda: rc.reduce_and_score(P, tabs, k=1), reps=5)
    print(f"K8 reduce64: {t*1e3:.3f} ms  {outs*8/t/1e9:.1f} GB/s", ts)
    P32 = P.float()
    t, ts = ev_time(lambda: rc.reduce_and_score(P32, tabs, k=1), reps=5)
    print(f"K8 reduce32: {t*1e3:.3f} ms  {outs*4/t/1e9:.1f} GB/s", ts)
# cuBLAS DGEMM reference peak
X = torch.rand(8192, 8192, dtype=torch.float64, device=dev); Y = torch.rand(8192, 8192, dtype=torch.float64, device=dev)
t, ts = ev_time(lambda: torch.matmul(X, Y), reps=3)
print(f"cuBLAS DGEMM 8192^3: {2*8192**3/t/1e12:.2f} TFLOP/s")
