"""Where the host time of mis.cut_cost goes at the C4-circuit shape (cProfile of a 4-set call after warm-up)."""
import cProfile, importlib, io, os, pstats, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
import bench
from qdc_b200 import mis
from qdc_b200.engine import B200Backend
dev = torch.device("cuda:0")
G, circ, frags, wpm = bench.c4_circuit_fragments()
used = circ.parameters
be = B200Backend(device=dev, graph=G)
os.environ["QDC_CONTRACTION"] = "i8:auto"
for n_sets in (1, 2, 4):
    table = np.random.default_rng(5).uniform(0, 2 * np.pi, (n_sets, len(used)))
    for it in range(4):
        torch.cuda.synchronize(); t = time.time()
        mis.cut_cost(frags, wpm, 1, be, params=table, param_order=used, mode="likely")
        torch.cuda.synchronize(); dt = (time.time() - t) * 1e3
    print(f"n_sets {n_sets}: {dt:.2f} ms per call, {dt / n_sets:.2f} per evaluation", flush=True)
pr = cProfile.Profile(); pr.enable()
mis.cut_cost(frags, wpm, 1, be, params=table, param_order=used, mode="likely")
torch.cuda.synchronize(); pr.disable()
st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(28); print(st.getvalue()[-4500:])
