"""Device timing of the int8-sliced contraction (qdc_contract_i8) against the FP64 DMMA kernel
on the bench operands (C4 shape by default), with the difference of the three sums."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import recombine as rc
import bench

dev = torch.device("cuda:0")
rb = int(sys.argv[1]) if len(sys.argv) > 1 else 17
cb = int(sys.argv[2]) if len(sys.argv) > 2 else 17
slice_list = [int(s) for s in sys.argv[3].split(",")] if len(sys.argv) > 3 else [6, 5, 4]
n = rb + cb
A, Bt = bench.make_operands(rb, cb)
A, Bt = torch.from_numpy(A).to(dev), torch.from_numpy(Bt).to(dev)
kdim = A.shape[1]
G, adj = bench.make_graph(n)
node_mask = (1 << n) - 1
tabs = rc.build_tables(rb, cb, bench.make_bitpos(n, cb), adj, node_mask, device=dev)
outs = float(1 << n)


def run(method, slices=6, reps=3):
    ts, res = [], None
    for it in range(reps + 1):
        evs = []
        res = rc.contract_and_score(A, Bt, tabs, scale=1.0 / 16, k=1, kernel_events=evs,
                                    method=method, slices=slices)
        torch.cuda.synchronize()
        if it:
            ts.append(evs[0][0].elapsed_time(evs[0][1]))
    return min(ts), res.get()


t64, r64 = run("f64")
print(f"f64 DMMA     : {t64:8.2f} ms  {outs / t64 / 1e6:8.2f} Gprobs/s  {2 * outs * kdim / t64 / 1e9:6.1f} TFLOP/s-equiv", r64)
for S in slice_list:
    t, r = run("i8", S)
    pairs = S * (S + 1) // 2
    print(f"i8 S={S} ({pairs:2d} pairs): {t:8.2f} ms  {outs / t / 1e6:8.2f} Gprobs/s  "
          f"{2 * outs * kdim * pairs / t / 1e9:7.1f} int8 TOP/s  "
          f"d_sum_p {r['sum_p'] - r64['sum_p']:.2e} d_pw_indset {r['sum_pw_indset'] - r64['sum_pw_indset']:.2e}")

