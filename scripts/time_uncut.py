"""Ad-hoc timing of the uncut DQVA cost (config C5 shape): statevector engine + amplitude scoring."""
import importlib, os, sys, time
import numpy as np, torch, networkx as nx
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import mis, recombine as rc
from qdc_b200.engine import B200Backend
from qdc_b200.ansatz import dqv_ansatz
from qdc_b200.circuit import Parameter


def case(n, n_sets, tile_bits, coalesce, fused=True):
    dev = torch.device("cuda:0")
    if n % 2 == 0:
        G = nx.random_regular_graph(3, n, seed=2)
    else:                                   # the bench's C5 graph: 4-regular on n + 1 nodes, one removed
        G = nx.random_regular_graph(4, n + 1, seed=2)
        G.remove_node(n)
    syms = [Parameter(f"var_{i}") for i in range(n + 1)]
    circ = dqv_ansatz.gen_dqva(G, P=1, params=syms, init_state="0" * n, barriers=0, fused=fused,
                               mixer_order=list(range(n)))
    be = B200Backend(device=dev, tile_bits=tile_bits, coalesce_bits=coalesce)
    used = circ.parameters
    prog, auto = be.compile(circ, param_order=used)
    table = np.random.default_rng(2).uniform(0, 2 * np.pi, (n_sets, len(used)))
    nq = prog.n_qubits
    states = torch.empty(n_sets, 1 << nq, dtype=torch.complex128, device=dev)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(n)] + [0] * (nq - n)
    tabs = rc.build_tables(nq - 14, 14, list(range(nq)), adj, (1 << n) - 1, device=dev, index_flip=prog.flip_mask)

    def ev(fn, reps=3):
        fn(); torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e-3)
        return min(ts)
    t_run = ev(lambda: be.run(prog, auto, table, out=states))
    t_red = ev(lambda: [rc.reduce_and_score(torch.view_as_real(states[s]).reshape(-1), tabs, threshold=2.0,
                                            cand_cap=1, amplitudes=True) for s in range(n_sets)])
    res = rc.reduce_and_score(torch.view_as_real(states[0]).reshape(-1), tabs, threshold=2.0, cand_cap=1,
                              amplitudes=True).get()
    state_gb = 16.0 * (1 << nq) / 1e9
    traffic = (2 * prog.n_groups - 1) * state_gb * n_sets          # first group writes only
    print(f"uncut DQVA n={n}(+anc={nq - n}) sets={n_sets} t={prog.tile_bits} c={coalesce} groups={prog.n_groups} "
          f"ops={prog.n_ops}: run {t_run*1e3:.2f} ms ({traffic/t_run:.0f} GB/s of group traffic), "
          f"score {t_red*1e3:.2f} ms ({state_gb*n_sets/t_red:.0f} GB/s); {n_sets/(t_run+t_red):.2f} evals/s; "
          f"sum_p={res['sum_p']:.12f} cost={-res['sum_pw']:.6f}", flush=True)


if __name__ == "__main__":
    args = [int(a) for a in sys.argv[1:]]
    n = args[0] if args else 26
    sets = args[1] if len(args) > 1 else 2
    for t, c in [(12, 4), (12, 5), (12, 6), (13, 5), (11, 4)]:
        case(n, sets, t, c)
