#!/bin/bash
mkdir -p gpurun_out
cat > /tmp/pair_one.py <<'P'
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
S = 5
pa = rc.prepare_i8(A, tabs.row_wok, slices=S)
pb = rc.prepare_i8(Bt, tabs.col_wok, tabs.col_x, slices=S, column_role=True)
for pair in ("0", "1"):
    os.environ["QDC_I8_PAIR"] = pair
    for it in range(2):
        r = rc.contract_and_score(pa, pb, tabs, scale=1.0 / 16, k=1, method="i8", slices=S)
        torch.cuda.synchronize()
print("done")
P
timeout 600 python /tmp/pair_one.py > gpurun_out/pair_one.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:contract_i8_kernel -s 1 -c 3 \
   -o gpurun_out/r2_i8_pair python /tmp/pair_one.py > gpurun_out/pair_ncu.log 2>&1
echo "ncu rc=$?"; tail -4 gpurun_out/pair_ncu.log
