"""Stage times of one avg_cost evaluation on the real cut-DQVA fragments of the C4 shape (bench.c4_circuit_fragments)."""
import importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
import bench
from qdc_b200.engine import B200Backend
from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm

dev = torch.device("cuda:0")
G, circ, frags, wpm = bench.c4_circuit_fragments()
used = circ.parameters
be = B200Backend(device=dev, graph=G)
table = np.random.default_rng(5).uniform(0, 2 * np.pi, (1, len(used)))
os.environ["QDC_CONTRACTION"] = sys.argv[1] if len(sys.argv) > 1 else "i8:auto"


def timed(fn):
    torch.cuda.synchronize(); t = time.time(); r = fn(); torch.cuda.synchronize(); return r, (time.time() - t) * 1e3


for it in range(3):
    data, t_collect = timed(lambda: qmm.collect_fragment_data(frags, wpm, shots=1, tomography_backend=be, params=table, param_order=used))
    direct, t_direct = timed(lambda: qmm.direct_fragment_model(data))
    ml, t_ml = timed(lambda: qmm.maximum_likelihood_model(direct))
    dist, t_rec = timed(lambda: qmm.recombine_fragment_models(ml, wpm, backend=be, store=False))
    print(f"run {it}: collect {t_collect:.2f} ms, direct model {t_direct:.2f}, ML {t_ml:.2f}, recombine+score {t_rec:.2f}; "
          f"tensors {[tuple(m.T.shape) for m in ml]}", flush=True)
for i, m in enumerate(ml):
    print("fragment", i, "ML info (needed full path, Newton steps, t0):", qmm.ml_info_host(m))
    _, t_one = timed(lambda: qmm.maximum_likelihood_model(direct[i]))
    _, t_forced = timed(lambda: qmm.maximum_likelihood_model(direct[i], force_full=True))
    print(f"   ML alone {t_one:.2f} ms, forced full path {t_forced:.2f} ms, blocks {m.T.shape[1]}")
d = dist[0] if isinstance(dist, list) else dist
print("score", d.score.get() if d.score is not None else None, "slices", getattr(d.score, "slices", None))
