"""Why the peaked-operand variant of the bench is slower than the flat one: same operands, candidate threshold on / off,
pair / single kernel, hot rows only / hot columns only."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import recombine as rc
import bench
dev = torch.device("cuda:0")
rb, cb = 17, 17
n = rb + cb
A, Bt = bench.make_operands(rb, cb)
A, Bt = torch.from_numpy(A).to(dev), torch.from_numpy(Bt).to(dev)
G, adj = bench.make_graph(n)
tabs = rc.build_tables(rb, cb, bench.make_bitpos(n, cb), adj, (1 << n) - 1, device=dev)
g = torch.Generator(device="cpu").manual_seed(11)
hr = torch.randperm(A.shape[0], generator=g)[:64].to(dev)
hc = torch.randperm(Bt.shape[0], generator=g)[:64].to(dev)


def run(Ax, Bx, thr, tag):
    ts = []
    for it in range(3):
        evs = []
        r = rc.contract_and_score(Ax, Bx, tabs, scale=1.0 / 16, k=1, threshold=thr, kernel_events=evs, method="i8", slices=5)
        torch.cuda.synchronize()
        if it: ts.append(evs[0][0].elapsed_time(evs[0][1]))
    print(f"{tag:44s} threshold {thr:g}: kernel {min(ts):7.2f} ms, candidates {r.get()['n_candidates']}", flush=True)


for pair in ("1", "0"):
    os.environ["QDC_I8_PAIR"] = pair
    Ap, Bp = A.clone(), Bt.clone(); Ap[hr] *= 8192.0; Bp[hc] *= 8192.0
    run(A, Bt, 1e-5, f"pair={pair} flat")
    run(Ap, Bp, 1e-5, f"pair={pair} hot rows x hot columns")
    run(Ap, Bp, 1.0, f"pair={pair} hot rows x hot columns")
    run(Ap, Bt, 1e-5, f"pair={pair} hot rows only")
    run(A, Bp, 1e-5, f"pair={pair} hot columns only")
