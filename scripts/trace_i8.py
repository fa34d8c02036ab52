"""Timeline of a few tiles of contract_i8_kernel in CTA 0 (diagnostic build: make EXTRA=-DQDC_I8_TRACE).
Prints, per traced warp, the events with clocks relative to the first event of tile 200."""
import ctypes as C, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import recombine as rc, _lib
import bench

lib = _lib.load()
dev = torch.device("cuda:0")
rb, cb = 17, 17
n = rb + cb
A, Bt = bench.make_operands(rb, cb)
A, Bt = torch.from_numpy(A).to(dev), torch.from_numpy(Bt).to(dev)
G, adj = bench.make_graph(n)
tabs = rc.build_tables(rb, cb, bench.make_bitpos(n, cb), adj, (1 << n) - 1, device=dev)
S = int(sys.argv[1]) if len(sys.argv) > 1 else 5
pa = rc.prepare_i8(A, tabs.row_wok, slices=S)
pb = rc.prepare_i8(Bt, tabs.col_wok, tabs.col_x, slices=S, column_role=True)
raw = C.CDLL(lib._name)
names = {0: "producer", 1: "issuer0", 2: "issuer1", 3: "issuer2", 4: "epi first", 5: "epi last"}
KIND = {1: "copy slice", 2: "b_full j", 3: "acc_empty d", 4: "issue pair", 5: "acc_full d", 6: "release d", 7: "epi tile"}
for pair in ("0", "1"):
    os.environ["QDC_I8_PAIR"] = pair
    ev = (C.c_ulonglong * (8 * 1024))()
    cnt = (C.c_int * 8)()
    rc.contract_and_score(pa, pb, tabs, scale=1.0 / 16, k=1, method="i8", slices=S)
    torch.cuda.synchronize()
    raw.qdc_i8_debug_trace(ev, cnt)                       # warm-up run discarded
    rc.contract_and_score(pa, pb, tabs, scale=1.0 / 16, k=1, method="i8", slices=S)
    torch.cuda.synchronize()
    raw.qdc_i8_debug_trace(ev, cnt)
    e = np.frombuffer(ev, dtype=np.uint64).reshape(8, 512, 2)
    allev = []
    for w in range(6):
        for k in range(min(cnt[w], 512)):
            code, t = int(e[w, k, 0]), int(e[w, k, 1])
            allev.append((t, w, code >> 32, code & 0xffff))
    allev.sort()
    t0 = allev[0][0]
    print(f"==== S={S} pair={pair}: {len(allev)} events, span {allev[-1][0] - t0} clk for 4 tiles")
    for t, w, tile, code in allev:
        kind, arg = code >> 8, code & 0xff
        txt = f"({arg >> 4},{arg & 15})" if kind == 4 else str(arg)
        print(f"{t - t0:8d}  tile {tile}  {names[w]:10s} {KIND[kind]:12s} {txt}")
