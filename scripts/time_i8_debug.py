"""Where the int8 contraction's time goes: the kernel at the C4 shape with the diagnostic knobs of
csrc/contract_i8.cu (QDC_I8_DEBUG: 1 = epilogue without arithmetic, 2 = streamed operand copied for
the first tile only, 4 = no MMAs) and tile width / ring depth overrides.  Results are wrong when a
knob is set; only the CUDA-event time of plan + contraction + finish is printed."""
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
pa = {S: rc.prepare_i8(A, tabs.row_wok, slices=S) for S in (5, 6)}
pb = {S: rc.prepare_i8(Bt, tabs.col_wok, tabs.col_x, slices=S, column_role=True) for S in (5, 6)}


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


cases = [(5, {}), (5, {"QDC_I8_DEBUG": "1"}), (5, {"QDC_I8_DEBUG": "2"}), (5, {"QDC_I8_DEBUG": "3"}),
         (5, {"QDC_I8_DEBUG": "4"}), (5, {"QDC_I8_DEBUG": "5"}), (5, {"QDC_I8_DEBUG": "7"}),
         (5, {"QDC_I8_TN": "64"}), (5, {"QDC_I8_TN": "64", "QDC_I8_DEBUG": "1"}), (5, {"QDC_I8_TN": "64", "QDC_I8_DEBUG": "3"}),
         (5, {"QDC_I8_TN": "64", "QDC_I8_TMEM_A": "0"}), (5, {"QDC_I8_TN": "64", "QDC_I8_STAGES": "2"}),
         (6, {}), (6, {"QDC_I8_DEBUG": "1"}), (6, {"QDC_I8_DEBUG": "3"}), (6, {"QDC_I8_DEBUG": "4"})]
for S, env in cases:
    print(f"S={S} {str(env):60s} {run(S, env):8.2f} ms", flush=True)
