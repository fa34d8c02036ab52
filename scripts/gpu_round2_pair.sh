#!/bin/bash
mkdir -p gpurun_out
echo "== tests with the CTA-pair kernel (default on)"
timeout 900 python -m pytest tests/test_gpu_contract_score.py -x -q -m gpu 2>&1 | tail -15
echo "== timing"
cat > /tmp/pair_time.py <<'P'
import importlib, os, sys
import torch
sys.path.insert(0, os.getcwd())
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
for S in (5, 4):
    pa = rc.prepare_i8(A, tabs.row_wok, slices=S)
    pb = rc.prepare_i8(Bt, tabs.col_wok, tabs.col_x, slices=S, column_role=True)
    out = {}
    for pair in ("0", "1"):
        os.environ["QDC_I8_PAIR"] = pair
        ts = []
        for it in range(4):
            evs = []
            r = rc.contract_and_score(pa, pb, tabs, scale=1.0 / 16, k=1, kernel_events=evs, method="i8", slices=S)
            torch.cuda.synchronize()
            if it: ts.append(evs[0][0].elapsed_time(evs[0][1]))
        out[pair] = (min(ts), r.get()["sum_pw_indset"])
        print(f"S={S} pair={pair}: {min(ts):8.2f} ms  cost {out[pair][1]!r}", flush=True)
    print("  cost difference pair - single:", out["1"][1] - out["0"][1])
P
timeout 600 python /tmp/pair_time.py 2>&1 | tail -12 | tee gpurun_out/pair_time.log
