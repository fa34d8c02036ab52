"""The int8 contraction kernel alone at the C4 shape (operands prepared once) for the tile widths and
slice counts in use: CUDA-event time of plan + contraction + finish.  (The diagnostic knobs this
script drove in round 2 -- epilogue without arithmetic, no copies, no MMAs -- are gone from the
kernel; their numbers are in profiles/r2_i8_debug_knobs.log.)"""
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
pa = {S: rc.prepare_i8(A, tabs.row_wok, slices=S) for S in (4, 5, 6)}
pb = {S: rc.prepare_i8(Bt, tabs.col_wok, tabs.col_x, slices=S, column_role=True) for S in (4, 5, 6)}


def run(S, env, reps=3):
    for k in ("QDC_I8_DEBUG", "QDC_I8_TN", "QDC_I8_STAGES", "QDC_I8_TMEM_A"):
        os.environ.pop(k, None)
    os.environ.update(env)
    ts = []
    for it in range(reps + 1):
        evs = []
        rc.contract_and_score(pa[S], pb[S], tabs, scale=1.0 / 16, k=1, kernel_events=evs, method="i8", slices=S)
        torch.cuda.synchronize()
        if it:
            ts.append(evs[0][0].elapsed_time(evs[0][1]))
    return min(ts)


cases = [(5, {"QDC_I8_TN": "128"}), (5, {"QDC_I8_TN": "96"}), (5, {"QDC_I8_TN": "64"}),
         (6, {}), (4, {"QDC_I8_TN": "128"}), (4, {"QDC_I8_TN": "64"})]
for S, env in cases:
    print(f"S={S} {str(env):40s} {run(S, env):8.2f} ms", flush=True)
